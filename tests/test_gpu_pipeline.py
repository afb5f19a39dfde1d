"""GPU parity of the batched pipeline handle (tk_b200_pipeline_*) and of the reference-typed
single-frame post-process, against the oracle and against the individual launchers."""
import numpy as np
import pytest
import torch

import synth
from oracle import oracle as O
from trackiellm_b200 import vision as V

pytestmark = pytest.mark.gpu

W, H = 320, 240           # small frames keep the oracle fast; full 1080p is covered in test_gpu_preprocess
P, NC = 2000, 80


def inputs(n, seed0=0, p=P):
    frames = np.stack([synth.rgb_frame(W, H, seed0 + i) for i in range(n)])
    logits = np.stack([synth.logits_tensor(seed0 + i, num_proposals=p, objects=12, dups=6, adversarial=(i == 0))
                       for i in range(n)])
    return frames, logits


def oracle_dets(logit, restore=None):
    c, _ = O.decode(logit, 0.5, cap=4096)
    d = O.nms_classwise(c, 0.5, 500)
    if restore:
        d = O.restore_coords(d, *restore)
    return d.view(V.CAND_DTYPE)


def make_pipe(n, **kw):
    args = dict(max_batch=n, chunk_frames=3, width=W, height=H, tensor_w=64, tensor_h=48, out_dtype=torch.float32,
                num_proposals=P, num_classes=NC, conf_thr=0.5, iou_thr=0.5, max_detections=500)
    args.update(kw)
    return V.Pipeline(**args)


def check_dets(dets, counts, logits, restore=None):
    lists = V.candidates_to_numpy(dets if dets.is_cuda else dets.cuda(), counts if counts.is_cuda else counts.cuda())
    for i, got in enumerate(lists):
        ref = oracle_dets(logits[i], restore)
        assert got.tobytes() == ref.tobytes(), f"frame {i}: {len(got)} vs {len(ref)} detections"


@pytest.mark.parametrize("w,h,fmt", [(320, 240, "rgb"), (250, 240, "rgb"), (100, 60, "rgb"), (320, 240, "i420"), (320, 240, "rgba")])
def test_process_host_frame_paths_bitexact(w, h, fmt):
    """Page-locked frames: sparse geometries with 16-byte rows are read IN PLACE over PCIe by the row ring, the others
    (row bytes not a multiple of 16, dense geometry) are copied in chunks — every path bit-exact, twice per handle."""
    n = 5
    rng = np.random.default_rng(w + h)
    if fmt == "i420":
        bufs = np.stack([synth.i420_frame(w, h, 40 + i) for i in range(n)])
        rgb = [O.i420_to_rgb8(b, w, h) for b in bufs]
        kw = dict(fmt=V.PIXEL_I420)
    elif fmt == "rgba":
        rgb = [rng.integers(0, 256, (h, w, 3), dtype=np.uint8) for _ in range(n)]
        bufs = np.stack([np.concatenate([r, rng.integers(0, 256, (h, w, 1), dtype=np.uint8)], axis=2) for r in rgb]).reshape(n, -1)
        kw = dict(fmt=V.PIXEL_RGBA8)
    else:
        rgb = [rng.integers(0, 256, (h, w, 3), dtype=np.uint8) for _ in range(n)]
        bufs = np.stack(rgb).reshape(n, -1)
        kw = dict(fmt=V.PIXEL_RGB8)
    with make_pipe(n, width=w, height=h, chunk_frames=2, arith=V.ARITH_EXACT, **kw) as pipe:
        hf = torch.from_numpy(bufs).pin_memory()
        ht = torch.empty((n, 3, 48, 64), dtype=torch.float32).pin_memory()
        for _ in range(2):
            ht.zero_()
            pipe.process_host(hf, None, None, None, tensor=ht)
            for i in range(n):
                ref = O.preprocess_rgb8(rgb[i], 64, 48)
                assert np.array_equal(ht[i].numpy().view(np.uint32), ref.view(np.uint32)), (i, w, h, fmt)


@pytest.mark.parametrize("mode", [0, 1, 2])
def test_process_host_logits_modes_full_head(mode):
    """25200-row heads (>= 65536 rows per batch: mode 0 packs the objectness column with host threads and decode reads
    only the surviving rows in place), every mode against the oracle; post-process only (no frames)."""
    n = 4
    logits = np.stack([synth.logits_tensor(70 + i, adversarial=(i == 1)) for i in range(n)])
    with make_pipe(n, num_proposals=synth.NUM_PROPOSALS, stage_host_logits=mode) as pipe:
        _, hd, hc = pipe.alloc_outputs(n, pinned=True)
        hl = torch.from_numpy(logits).pin_memory()
        for _ in range(2):
            hd.zero_(); hc.zero_()
            pipe.process_host(None, hl, hd, hc)
            check_dets(hd, hc, logits)


def test_process_device_vs_oracle():
    n = 7
    frames, logits = inputs(n)
    with make_pipe(n) as pipe:
        t, d, c = pipe.alloc_outputs(n)
        pipe.process_device(torch.from_numpy(frames).cuda(), torch.from_numpy(logits).cuda(), t, d, c)
        torch.cuda.synchronize()
        for i in range(n):
            ref = O.preprocess_rgb8(frames[i], 64, 48)
            assert np.array_equal(t[i].cpu().numpy().view(np.uint32), ref.view(np.uint32))
        check_dets(d, c, logits)


@pytest.mark.parametrize("stage", [0, 1, 2])     # objectness column packed on the host / tensor staged through HBM / read in place over PCIe
@pytest.mark.parametrize("n,chunk", [(1, 1), (7, 3), (8, 8), (10, 4), (5, 16)])
def test_process_host_matches_oracle_and_device(n, chunk, stage):
    frames, logits = inputs(n, seed0=10)
    with make_pipe(n, chunk_frames=chunk, out_dtype=torch.float16, stage_host_logits=stage) as pipe:
        _, hd, hc = pipe.alloc_outputs(n, pinned=True)
        ht = torch.empty((n, 3, 48, 64), dtype=torch.float16).pin_memory()
        hf = torch.from_numpy(frames).pin_memory()
        hl = torch.from_numpy(logits).pin_memory()
        for _ in range(2):                                   # twice: slot reuse across calls
            hd.zero_(); hc.zero_(); ht.zero_()
            pipe.process_host(hf, hl, hd, hc, tensor=ht)
            check_dets(hd, hc, logits)
            for i in range(n):
                ref = O.f32_to_f16(O.preprocess_rgb8(frames[i], 64, 48)).astype(np.float32)
                got = ht[i].numpy().astype(np.float32)
                assert np.all(np.abs(got - ref) <= 1e-2 + 1e-2 * np.abs(ref))
        # pageable (unpinned) host memory also works
        hd2 = torch.zeros_like(hd); hc2 = torch.zeros_like(hc)
        pipe.process_host(frames, logits, hd2.numpy(), hc2.numpy())
        assert torch.equal(hd2[:, :1], hd[:, :1]) and torch.equal(hc2, hc)
        # logits already on the device
        hd3 = torch.zeros_like(hd).pin_memory(); hc3 = torch.zeros_like(hc).pin_memory()
        pipe.process_host(hf, torch.from_numpy(logits).cuda(), hd3, hc3, logits_on_device=True)
        check_dets(hd3, hc3, logits)


def test_pre_only_and_post_only():
    n = 4
    frames, logits = inputs(n, seed0=20)
    with make_pipe(n) as pipe:
        ht = torch.zeros((n, 3, 48, 64), dtype=torch.float32).pin_memory()
        pipe.process_host(torch.from_numpy(frames).pin_memory(), None, None, None, tensor=ht)
        for i in range(n):
            assert np.array_equal(ht[i].numpy().view(np.uint32), O.preprocess_rgb8(frames[i], 64, 48).view(np.uint32))
        _, hd, hc = pipe.alloc_outputs(n, pinned=True)
        pipe.process_host(None, torch.from_numpy(logits).pin_memory(), hd, hc)
        check_dets(hd, hc, logits)


def test_restore_coordinates_and_threshold_update():
    n = 3
    frames, logits = inputs(n, seed0=30)
    with make_pipe(n, restore_coordinates=True) as pipe:
        t, d, c = pipe.alloc_outputs(n)
        pipe.process_device(torch.from_numpy(frames).cuda(), torch.from_numpy(logits).cuda(), t, d, c)
        torch.cuda.synchronize()
        check_dets(d, c, logits, restore=(W, H, 64, 48))
        pipe.update_thresholds(0.7, 0.3)                      # tk_object_detector_update_thresholds
        pipe.process_device(None, torch.from_numpy(logits).cuda(), None, d, c)
        torch.cuda.synchronize()
        for i, got in enumerate(V.candidates_to_numpy(d, c)):
            cands, _ = O.decode(logits[i], 0.7, cap=4096)
            ref = O.restore_coords(O.nms_classwise(cands, 0.3, 500), W, H, 64, 48).view(V.CAND_DTYPE)
            assert got.tobytes() == ref.tobytes()


def test_stage_timing_and_launch_count():
    import ctypes as C
    from trackiellm_b200 import _capi as capi
    n = 4
    frames, logits = inputs(n, seed0=40)
    lib = capi.lib()
    lib.tk_b200_launch_count.restype = C.c_ulonglong
    with make_pipe(n) as pipe:
        t, d, c = pipe.alloc_outputs(n)
        df, dl = torch.from_numpy(frames).cuda(), torch.from_numpy(logits).cuda()
        pipe.stage_timing(5)
        before = lib.tk_b200_launch_count()
        for _ in range(5):
            pipe.process_device(df, dl, t, d, c)
        torch.cuda.synchronize()
        assert lib.tk_b200_launch_count() - before == 15      # pre + decode + nms per call
        ms, calls = pipe.stage_times()
        assert calls == 5 and all(m > 0 for m in ms)
    for mode in (2, 3):                                       # the fused post-process forms: two launches per call
        with make_pipe(n, fuse_postprocess=mode) as pipe:
            t, d, c = pipe.alloc_outputs(n)
            pipe.stage_timing(5)
            before = lib.tk_b200_launch_count()
            for _ in range(5):
                pipe.process_device(df, dl, t, d, c)
            torch.cuda.synchronize()
            assert lib.tk_b200_launch_count() - before == 10
            ms, calls = pipe.stage_times()
            assert calls == 5 and ms[0] > 0 and ms[1] > 0 and ms[2] == 0


def test_errors():
    with make_pipe(2) as pipe:
        t, d, c = pipe.alloc_outputs(4)
        fr = torch.zeros((4, H, W, 3), dtype=torch.uint8, device="cuda")
        with pytest.raises(V.capi.TkError):                   # more frames than max_batch
            pipe.process_device(fr, None, t, d, c)
    with pytest.raises(V.capi.TkError):
        make_pipe(0)
    with pytest.raises(V.capi.TkError):
        make_pipe(1, max_candidates=5000)


def test_postprocess_detections_reference_result_type():
    logit = synth.logits_tensor(5, adversarial=True)
    labels = [f"class{i}" for i in range(80)]
    res = V.postprocess_detections(logit, conf_thr=0.5, iou_thr=0.5, max_detections=500, class_labels=labels,
                                   frame_wh=(1920, 1080), input_wh=(640, 640))
    ref = oracle_dets(logit, restore=(1920, 1080, 640, 640))
    assert len(res) == len(ref)
    for (cid, lab, conf, box), r in zip(res, ref):
        assert cid == r["class_id"] and lab == f"class{cid}"
        assert np.float32(conf).tobytes() == np.float32(r["confidence"]).tobytes() or (conf != conf and r["confidence"] != r["confidence"])
        assert box == (r["x"], r["y"], r["w"], r["h"])
    # nothing above the threshold -> empty, NULL results
    assert V.postprocess_detections(logit, conf_thr=2.0) == []


def test_shard_harness_one_process_several_pipelines():
    """tk_b200_shard_*: one process, one host thread + one pipeline per entry of device_ids, contiguous frame blocks,
    outputs written straight into the job's host arrays.  On a single-GPU box the ids repeat device 0 (three pipelines,
    three host threads, one device); with more GPUs every device takes a block.  Against one pipeline over the whole
    batch and against the oracle, twice per handle, with a threshold update in between."""
    n_dev = max(torch.cuda.device_count(), 1)
    ids = list(range(n_dev)) if n_dev > 1 else [0, 0, 0]
    n = 11                                                   # not a multiple of the device count: ragged blocks
    frames, logits = inputs(n, seed0=300)
    kw = dict(max_batch=n, chunk_frames=2, width=W, height=H, tensor_w=64, tensor_h=48, out_dtype=torch.float32,
              arith=V.ARITH_EXACT, num_proposals=P, num_classes=NC, conf_thr=0.5, iou_thr=0.5, max_detections=500)
    hf, hl = torch.from_numpy(frames).pin_memory(), torch.from_numpy(logits).pin_memory()
    with V.Shard(ids, **kw) as sh:
        hd = torch.zeros((n, 500, 7), dtype=torch.int32).pin_memory()
        hc = torch.zeros((n,), dtype=torch.int32).pin_memory()
        ht = torch.zeros((n, 3, 48, 64), dtype=torch.float32).pin_memory()
        for rep in range(2):
            hd.zero_(); hc.zero_(); ht.zero_()
            sh.process_host(hf, hl, hd, hc, tensor=ht)
            check_dets(hd, hc, logits)
            for i in range(n):
                assert np.array_equal(ht[i].numpy().view(np.uint32), O.preprocess_rgb8(frames[i], 64, 48).view(np.uint32))
        flat, offs = sh.gather(hd, hc)
        assert int(offs[-1]) == int(hc.sum()) and len(offs) == n + 1
        for f in range(n):
            assert np.array_equal(flat[int(offs[f]):int(offs[f + 1])], hd[f, :int(hc[f])].numpy())
        sh.update_thresholds(0.9, 0.5)                       # fewer rows pass on every device
        sh.process_host(None, hl, hd, hc)
        for i in range(n):
            c, _ = O.decode(logits[i], 0.9, cap=4096)
            ref = O.nms_classwise(c, 0.5, 500).view(V.CAND_DTYPE)
            assert hd[i, :int(hc[i])].numpy().tobytes() == ref.tobytes()


def test_shard_harness_from_c(tmp_path):
    """tests/c/tk_shard_test.c: ONE C process drives every visible device (three pipelines on device 0 when there is
    one) through tk_b200_shard_process_host and compares tensors, counts, detections and the gathered flat list with a
    single tk_b200_pipeline over the whole batch, byte for byte; the detection total is checked against the oracle."""
    import os
    import subprocess
    exe = os.path.join(os.path.dirname(__file__), "c", "tk_shard_test")
    if not os.path.exists(exe):
        pytest.fail("tests/c/tk_shard_test is not built: run __graft_entry__.build()")
    n = 13
    frames, logits = inputs(n, seed0=500)
    fp, lp = tmp_path / "frames.u8", tmp_path / "logits.f32"
    frames.tofile(fp)
    logits.tofile(lp)
    out = subprocess.run([exe, str(fp), str(lp), str(n), str(W), str(H), str(P), str(NC)], capture_output=True, text=True)
    assert out.returncode == 0, out.stdout + out.stderr
    words = out.stdout.split()
    total = int(words[words.index("detections") + 1])
    want = 0
    for i in range(n):
        c, _ = O.decode(logits[i], 0.5, cap=4096)
        want += len(O.nms_classwise(c, 0.5, 300))
    assert total == want > 0
    print(out.stdout.strip())


@pytest.mark.parametrize("act", [V.ACT_RAW, V.ACT_SIGMOID, V.ACT_SOFTMAX])
@pytest.mark.parametrize("p,n", [(31, 3), (2000, 5), (25200, 2), (4097, 70)])
def test_scan_inside_preprocess_equals_plain_decode(act, p, n, monkeypatch):
    """TK_B200_SCAN=1: process_device runs decode's pass 1 (objectness threshold -> ballot words) in an extra warp of every
    pre-process CTA and decode_kernel loads the words (measured slower, off by default).  Same detections, counts and
    tensors either way and against the oracle — ragged proposal counts (words that straddle nothing: per-frame bitmaps),
    more frames than fit one CTA's slice, every activation."""
    frames = np.stack([synth.rgb_frame(W, H, 900 + i) for i in range(min(n, 6))])
    frames = frames[np.arange(n) % len(frames)]
    logits = np.stack([synth.logits_tensor(900 + i, num_proposals=p, objects=10, dups=5, adversarial=(i == 0 and p >= 64),
                                           as_logit=(act != V.ACT_RAW)) for i in range(min(n, 6))])
    logits = logits[np.arange(n) % len(logits)]
    outs = []
    for scan in ("1", "0"):
        monkeypatch.setenv("TK_B200_SCAN", scan)
        with make_pipe(n, num_proposals=p, act=act) as pipe:
            dt, dd, dc = pipe.alloc_outputs(n)
            df, dl = torch.from_numpy(frames).cuda(), torch.from_numpy(logits).cuda()
            for _ in range(2):
                pipe.process_device(df, dl, dt, dd, dc)
            torch.cuda.synchronize()
            outs.append((dt.cpu().numpy().copy(), V.candidates_to_numpy(dd, dc)))
    assert np.array_equal(outs[0][0].view(np.uint32), outs[1][0].view(np.uint32))
    for i in range(n):
        assert outs[0][1][i].tobytes() == outs[1][1][i].tobytes(), i
    for i in range(min(n, 6)):
        c, _ = O.decode(logits[i], 0.5, act, cap=4096)
        assert outs[0][1][i].tobytes() == O.nms_classwise(c, 0.5, 500).view(V.CAND_DTYPE).tobytes()
