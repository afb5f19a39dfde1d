"""GPU parity: fused pre-process kernel (through the C-ABI) vs the oracle / golden vectors."""
import hashlib
import os

import numpy as np
import pytest
import torch

import synth
from oracle import oracle as O
from trackiellm_b200 import vision as V

pytestmark = pytest.mark.gpu
G = np.load(os.path.join(os.path.dirname(__file__), "golden", "golden.npz"))
ATOL, RTOL = 1e-5, 1e-5          # fp32 bar of the north star: |a-b| <= atol + rtol*|ref|
F16_TOL = 1e-2                   # fp16 bar of the north star


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def run(frames, ow, oh, **kw):
    d = torch.from_numpy(np.ascontiguousarray(frames)).cuda()
    out = V.preprocess_batch(d, ow, oh, **kw)
    torch.cuda.synchronize()
    return out.cpu().numpy()


@pytest.mark.parametrize("name,ow,oh", [("pre_small_a", 40, 40), ("pre_small_b", 20, 12), ("pre_small_ramp", 32, 32)])
def test_golden_small_bitexact(name, ow, oh):
    got = run(G[name + "_in"][None], ow, oh)[0]
    assert np.array_equal(bits(got), bits(G[name + "_out"]))


@pytest.mark.parametrize("name,w,h,seed,kind", [
    ("pre_cfg1_zeros", 640, 480, 0, "zeros"), ("pre_cfg1_random", 640, 480, 0, "random"),
    ("pre_cfg1_ramp", 640, 480, 0, "ramp"), ("pre_cfg1_white", 640, 480, 0, "white"),
    ("pre_1080p_random", 1920, 1080, 7, "random")])
def test_golden_sha_reference_entry_points(name, w, h, seed, kind):
    fr = synth.rgb_frame(w, h, seed, kind)
    want = G[name + "_sha256"].tobytes()
    # (1) the reference's host API, (2) the reference's GPU launcher, (3) the batched kernel
    a = V.preprocessor_resize_and_normalize_to_chw(fr, 640, 640)
    b = V.kernels_preprocess_image(torch.from_numpy(fr).cuda(), 640, 640).cpu().numpy()
    c = run(fr[None], 640, 640)[0]
    for got in (a, b, c):
        assert hashlib.sha256(np.ascontiguousarray(got).tobytes()).digest() == want


@pytest.mark.parametrize("w,h,ow,oh", [(640, 480, 640, 640), (1920, 1080, 640, 640), (333, 257, 640, 640), (64, 48, 640, 640),
                                       (1920, 1080, 256, 256), (1920, 1080, 641, 333), (2, 2, 9, 5), (3840, 2160, 640, 640),
                                       (640, 640, 640, 640), (100, 1000, 37, 1)])
@pytest.mark.parametrize("arith", [V.ARITH_EXACT, V.ARITH_FAST])
def test_fp32_vs_oracle(w, h, ow, oh, arith):
    fr = np.stack([synth.rgb_frame(w, h, s) for s in (0, 1)])
    got = run(fr, ow, oh, arith=arith)
    for i in range(2):
        ref = O.preprocess_rgb8(fr[i], ow, oh)
        if arith == V.ARITH_EXACT:
            assert np.array_equal(bits(got[i]), bits(ref))
        else:
            assert np.all(np.abs(got[i] - ref) <= ATOL + RTOL * np.abs(ref))


@pytest.mark.parametrize("arith", [V.ARITH_EXACT, V.ARITH_FAST, V.ARITH_HALF])
def test_fp16_vs_oracle(arith):
    fr = np.stack([synth.rgb_frame(1920, 1080, s) for s in range(3)])
    got = run(fr, 640, 640, out_dtype=torch.float16, arith=arith)
    for i in range(3):
        ref32 = O.preprocess_rgb8(fr[i], 640, 640)
        ref16 = O.f32_to_f16(ref32)
        if arith == V.ARITH_EXACT:
            assert np.array_equal(got[i].view(np.uint16), ref16.view(np.uint16))
        g = got[i].astype(np.float32)
        err = np.abs(g - ref32)
        print(f"arith {arith} frame {i}: max abs err {err.max():.3e}, max err/tol {np.max(err / (F16_TOL + F16_TOL * np.abs(ref32))):.3f}")
        assert np.all(err <= F16_TOL + F16_TOL * np.abs(ref32))


@pytest.mark.parametrize("kind", ["random", "ramp", "white", "zeros"])
@pytest.mark.parametrize("w,h,ow,oh,fmt", [(1920, 1080, 640, 640, "rgb"), (3840, 2160, 640, 640, "rgb"), (640, 480, 640, 640, "rgb"),
                                           (333, 257, 96, 80, "rgb"), (320, 200, 96, 96, "rgba")])
def test_half_arithmetic_within_fp16_bar(kind, w, h, ow, oh, fmt):
    """ARITH_HALF (half2 lerp) against the fp32 oracle at the north star's fp16 bar, incl. letterbox."""
    fr = synth.rgb_frame(w, h, 11, kind)
    src = np.concatenate([fr, np.full((h, w, 1), 9, np.uint8)], axis=2) if fmt == "rgba" else fr
    kw = dict(fmt=V.PIXEL_RGBA8) if fmt == "rgba" else {}
    for resize in (V.RESIZE_STRETCH, V.RESIZE_LETTERBOX):
        got = run(src[None], ow, oh, out_dtype=torch.float16, arith=V.ARITH_HALF, resize=resize, **kw)[0].astype(np.float32)
        ref = O.preprocess_rgb8(fr, ow, oh, resize=resize, pad_value=114.0)
        assert np.all(np.abs(got - ref) <= F16_TOL + F16_TOL * np.abs(ref))


def parity_patterns(w, h, value_pairs):
    """Frames whose pixel value depends only on (x parity, y parity): every 2x2 neighbourhood is one of the 16 corner
    combinations of {a, b}, and over the 16 tables every output pixel — i.e. every (xd, yd) the geometry produces —
    sees all 16 of them (each channel walks the tables with its own offset)."""
    yy, xx = np.meshgrid(np.arange(h), np.arange(w), indexing="ij")
    sel = ((yy & 1) * 2 + (xx & 1)).astype(np.uint8)
    for (a, b) in value_pairs:
        for k in range(16):
            fr = np.empty((h, w, 3), np.uint8)
            for c in range(3):
                fr[..., c] = np.where((((k + 5 * c) % 16) >> sel) & 1, b, a)
            yield fr


@pytest.mark.parametrize("w,h", [(1920, 1080), (640, 480), (3840, 2160)])
def test_half_arithmetic_adversarial_bound(w, h):
    """ARITH_HALF under the content that maximises its error terms: 0/255 checkerboards and every other
    corner combination of {0, 255}, {0, 128}, {127, 128}, {1, 254} at every (x-phase, y-phase) of the map — the xd / yd
    rounding terms scale with the corner differences, the fp16 rounding of the intermediates with their magnitude —
    plus single-pixel impulses.  The bar is 1e-2 + 1e-2|ref|; the assertion is HALF of it."""
    pairs = [(0, 255), (0, 128), (127, 128), (1, 254)]
    if w == 3840:
        pairs = pairs[:1]                                   # 4K: the 16 tables of {0, 255} (the oracle takes ~1 s per frame)
    worst = 0.0
    frames = list(parity_patterns(w, h, pairs))
    imp = np.zeros((h, w, 3), np.uint8)
    imp[::3, ::3] = 255                                     # impulses: isolated 255s, every x / y phase of the 3-pixel grid
    frames += [imp, 255 - imp]
    for i in range(0, len(frames), 16):
        chunk = np.stack(frames[i:i + 16])
        got = run(chunk, 640, 640, out_dtype=torch.float16, arith=V.ARITH_HALF).astype(np.float32)
        for j in range(len(chunk)):
            ref = O.preprocess_rgb8(chunk[j], 640, 640)
            worst = max(worst, float(np.max(np.abs(got[j] - ref) / (F16_TOL + F16_TOL * np.abs(ref)))))
    print(f"{w}x{h}: worst err / bar over {len(frames)} adversarial frames = {worst:.3f}")
    assert worst <= 0.5


def test_batch_equals_singles_and_is_deterministic():
    fr = np.stack([synth.rgb_frame(1920, 1080, s) for s in range(5)])
    a = run(fr, 640, 640)
    b = run(fr, 640, 640)
    assert np.array_equal(bits(a), bits(b))
    for i in range(5):
        assert np.array_equal(bits(a[i]), bits(run(fr[i:i + 1], 640, 640)[0]))


def test_stride_and_unaligned_fallback():
    w, h = 301, 77                       # row = 903 B: not 16-byte aligned -> cooperative copy path
    fr = synth.rgb_frame(w, h, 4)
    ref = O.preprocess_rgb8(fr, 128, 64)
    assert np.array_equal(bits(run(fr[None], 128, 64)[0]), bits(ref))
    stride = 1024                        # padded rows, TMA-eligible
    buf = np.zeros((1, h * stride), np.uint8)
    buf[0].reshape(h, stride)[:, : w * 3] = fr.reshape(h, w * 3)
    got = run(buf, 128, 64, width=w, height=h, stride=stride)[0]
    assert np.array_equal(bits(got), bits(ref))


def test_rgba8():
    fr = synth.rgb_frame(320, 200, 6)
    rgba = np.concatenate([fr, np.full((200, 320, 1), 77, np.uint8)], axis=2)
    got = run(rgba[None], 96, 96, fmt=V.PIXEL_RGBA8)[0]
    assert np.array_equal(bits(got), bits(O.preprocess_rgb8(fr, 96, 96)))


@pytest.mark.parametrize("w,h,ow,oh", [(1920, 1080, 640, 640), (321, 203, 96, 80)])
def test_bgr8_ffi_container_rows(w, h, ow, oh):
    """TK_VISION_FMT_BGR24 with rows aligned up to TK_SIMD_ALIGNMENT = 32 bytes
    (src/ffi/c_api/tk_ffi_api.c:878-892): tensor planes stay R, G, B, so the result is the
    reference pre-processor's on the channel-reversed frame, bit for bit."""
    fr = synth.rgb_frame(w, h, 9)
    stride = (3 * w + 31) // 32 * 32
    buf = np.full((h, stride), 0xAB, np.uint8)
    buf[:, :3 * w] = fr[:, :, ::-1].reshape(h, 3 * w)
    ref = O.preprocess_rgb8(fr, ow, oh)
    got = run(buf.reshape(1, -1), ow, oh, width=w, height=h, fmt=V.PIXEL_BGR8, stride=stride)[0]
    assert np.array_equal(bits(got), bits(ref))
    got16 = run(buf.reshape(1, -1), ow, oh, width=w, height=h, fmt=V.PIXEL_BGR8, stride=stride,
                out_dtype=torch.float16, arith=V.ARITH_HALF)[0].astype(np.float32)
    assert np.all(np.abs(got16 - ref) <= 1e-2 + 1e-2 * np.abs(ref))


@pytest.mark.parametrize("w,h", [(3840, 2160), (640, 480), (130, 66)])
def test_i420(w, h):
    buf = synth.i420_frame(w, h, 8)
    rgb = O.i420_to_rgb8(buf, w, h)
    got = run(buf[None], 640, 640, width=w, height=h, fmt=V.PIXEL_I420)[0]
    assert np.array_equal(bits(got), bits(O.preprocess_rgb8(rgb, 640, 640)))
    got16 = run(buf[None], 640, 640, width=w, height=h, fmt=V.PIXEL_I420, out_dtype=torch.float16, arith=V.ARITH_EXACT)[0]
    assert np.array_equal(got16.view(np.uint16), O.f32_to_f16(O.preprocess_rgb8(rgb, 640, 640)).view(np.uint16))


@pytest.mark.parametrize("w,h", [(640, 480), (1920, 1080), (480, 640)])
def test_letterbox(w, h):
    fr = synth.rgb_frame(w, h, 2)
    got = run(fr[None], 640, 640, resize=V.RESIZE_LETTERBOX, pad_value=114.0)[0]
    ref = O.preprocess_rgb8(fr, 640, 640, resize=O.RESIZE_LETTERBOX, pad_value=114.0)
    assert np.array_equal(bits(got), bits(ref))


def test_error_codes():
    from trackiellm_b200 import _capi as capi
    d = torch.zeros((1, 1, 8, 3), dtype=torch.uint8, device="cuda")
    with pytest.raises(capi.TkError) as e:
        V.preprocess_batch(d, 8, 8)
    assert e.value.code == capi.TK_ERROR_INVALID_DIMENSIONS
    with pytest.raises(capi.TkError):
        V.preprocess_batch(torch.zeros((1, 8, 8, 3), dtype=torch.uint8, device="cuda"), 8, 8, std=(0.0, 1.0, 1.0))


def test_full_size_property_constant_frames():
    # size-independent property at BASELINE cfg-2 size: a constant frame maps to a constant tensor
    fr = np.full((4, 1080, 1920, 3), 0, np.uint8)
    for i, v in enumerate((0, 1, 128, 255)):
        fr[i] = v
    got = run(fr, 640, 640)
    for i, v in enumerate((0, 1, 128, 255)):
        for c in range(3):
            want = ((np.float32(v) * np.float32(1.0 / 255.0)) - np.float32(O.IMAGENET_MEAN[c])) / np.float32(O.IMAGENET_STD[c])
            plane = got[i, c]
            assert np.all(np.abs(plane - want) <= ATOL + RTOL * abs(want))


@pytest.mark.parametrize("w,h,ow,oh,fmt", [(3840, 2160, 640, 640, "rgb"), (3840, 2160, 640, 640, "i420"),
                                           (1920, 1080, 640, 640, "rgb"), (1280, 720, 352, 352, "rgb"),
                                           (1000, 2000, 64, 24, "rgb")])
def test_ring_stress_many_rows_per_cta(w, h, ow, oh, fmt):
    """Many frames per launch -> many ring wrap-arounds per CTA (a 4K row that no output row samples
    used to be re-armed while still in flight); repeated launches must be identical, and match the oracle."""
    n = 12
    if fmt == "i420":
        fr = np.stack([synth.i420_frame(w, h, 100 + i) for i in range(n)])
        kw = dict(width=w, height=h, fmt=V.PIXEL_I420)
        ref_of = lambda i: O.preprocess_rgb8(O.i420_to_rgb8(fr[i], w, h), ow, oh)   # noqa: E731
    else:
        fr = np.stack([synth.rgb_frame(w, h, 100 + i) for i in range(n)])
        kw = {}
        ref_of = lambda i: O.preprocess_rgb8(fr[i], ow, oh)                         # noqa: E731
    d = torch.from_numpy(fr).cuda()
    first = None
    for _ in range(6):
        out = V.preprocess_batch(d, ow, oh, **kw)
        torch.cuda.synchronize()
        if first is None:
            first = out.clone()
        else:
            assert torch.equal(out, first)
    got = first.cpu().numpy()
    for i in (0, n // 2, n - 1):
        assert np.array_equal(bits(got[i]), bits(ref_of(i)))


def test_random_geometries_all_formats_bitexact():
    """60 seeded random (source, target, batch, format, stride) combinations — odd widths, up- and
    down-scaling in either axis, unaligned strides (no-TMA fallback), sparse and dense row staging —
    every output bit against the oracle (= the compiled reference arithmetic)."""
    rng = np.random.default_rng(2024)
    for case in range(60):
        fmt = ("rgb", "bgr", "rgba", "i420")[case % 4]
        w, h = int(rng.integers(2, 700)), int(rng.integers(2, 500))
        if fmt == "i420":
            w, h = max(4, w & ~1), max(4, h & ~1)
        ow, oh = int(rng.integers(1, 300)), int(rng.integers(1, 300))
        n = int(rng.integers(1, 5))
        bpp = {"rgb": 3, "bgr": 3, "rgba": 4, "i420": 1}[fmt]
        pad = int(rng.choice([0, 0, 1, 5, 16, 32])) if fmt != "i420" else 0
        stride = w * bpp + pad
        rgb = [rng.integers(0, 256, (h, w, 3), dtype=np.uint8) for _ in range(n)]
        if fmt == "i420":
            bufs = [synth.i420_frame(w, h, 7000 + 10 * case + i) for i in range(n)]
            rgb = [O.i420_to_rgb8(b, w, h) for b in bufs]
            flat = np.stack(bufs)
            kw = dict(width=w, height=h, fmt=V.PIXEL_I420)
        else:
            flat = np.full((n, h, stride), 0x5A, np.uint8)
            for i in range(n):
                px = rgb[i] if fmt != "bgr" else rgb[i][:, :, ::-1]
                if fmt == "rgba":
                    px = np.concatenate([px, rng.integers(0, 256, (h, w, 1), dtype=np.uint8)], axis=2)
                flat[i, :, : w * bpp] = px.reshape(h, w * bpp)
            flat = flat.reshape(n, h * stride)
            kw = dict(width=w, height=h, stride=stride,
                      fmt={"rgb": V.PIXEL_RGB8, "bgr": V.PIXEL_BGR8, "rgba": V.PIXEL_RGBA8}[fmt])
        got = run(flat, ow, oh, **kw)
        for i in range(n):
            ref = O.preprocess_rgb8(rgb[i], ow, oh)
            assert np.array_equal(bits(got[i]), bits(ref)), (case, fmt, w, h, ow, oh, n, stride, i)


def test_host_entry_point_threads_do_not_leak_device_memory():
    """The host-pointer entry points keep a stream and grow-only device buffers per calling thread (they were
    never released).  30 short-lived threads, each one 1080p call: the device's free memory after they are gone is what
    it was before them (within one allocator granule), and tk_b200_thread_release() frees the caller's own."""
    import threading
    from trackiellm_b200 import _capi as capi
    fr = synth.rgb_frame(1920, 1080, 1)
    ref = O.preprocess_rgb8(fr, 640, 640)
    V.preprocessor_resize_and_normalize_to_chw(fr, 640, 640)          # this thread's scratch exists from here on
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    errs = []

    def work():
        try:
            got = V.preprocessor_resize_and_normalize_to_chw(fr, 640, 640)
            if not np.array_equal(bits(got), bits(ref)):
                errs.append("mismatch")
        except Exception as e:      # noqa: BLE001
            errs.append(repr(e))

    for _ in range(30):
        t = threading.Thread(target=work)
        t.start(); t.join()
    assert not errs, errs
    import time
    for _ in range(40):              # join() returns when the Python function ends; the OS thread's thread_local destructors follow
        torch.cuda.synchronize()
        free1, _ = torch.cuda.mem_get_info()
        if free0 - free1 < 4 << 20:
            break
        time.sleep(0.05)
    assert free0 - free1 < 4 << 20, (free0, free1)       # 30 leaked scratches would be ~440 MB
    capi.lib().tk_b200_thread_release()
    torch.cuda.synchronize()
    free2, _ = torch.cuda.mem_get_info()
    assert free2 >= free1 + (8 << 20)                     # this thread's ~11 MB of scratch went back
    assert np.array_equal(bits(V.preprocessor_resize_and_normalize_to_chw(fr, 640, 640)), bits(ref))   # and it comes back on demand


_DYN_CHILD = r"""
import sys
import numpy as np, torch
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
import synth
from oracle import oracle as O
from trackiellm_b200 import vision as V
bits = lambda a: np.ascontiguousarray(a, np.float32).view(np.uint32)
for (w, h, ow, oh, n, fmt) in ((320, 200, 256, 192, 40, "rgb"), (1920, 1080, 640, 640, 24, "rgb"), (3840, 2160, 640, 640, 10, "i420")):
    if fmt == "i420":
        fr = np.stack([synth.i420_frame(w, h, 300 + i) for i in range(n)])
        kw = dict(width=w, height=h, fmt=V.PIXEL_I420)
        ref = lambda i: O.preprocess_rgb8(O.i420_to_rgb8(fr[i], w, h), ow, oh)
    else:
        fr = np.stack([synth.rgb_frame(w, h, 300 + i) for i in range(n)])
        kw = {}
        ref = lambda i: O.preprocess_rgb8(fr[i], ow, oh)
    d = torch.from_numpy(fr).cuda()
    outs = [V.preprocess_batch(d, ow, oh, **kw).cpu().numpy() for _ in range(4)]
    for o in outs[1:]:
        assert np.array_equal(bits(o), bits(outs[0])), "launches differ"
    for i in (0, n // 3, n - 1):
        assert np.array_equal(bits(outs[0][i]), bits(ref(i))), (w, h, i)
print("DYN-OK")
"""


@pytest.mark.parametrize("dyn", ["0", "100,1", "100,4", "55,3"])
def test_dynamic_row_share_extremes_bitexact(dyn):
    """The pre-process kernel hands part of the output rows out dynamically (PreK::static_rows, TK_B200_PRE_DYN).  The
    split is read once per process, so each setting runs in a child: all static, ALL rows dynamic in one-row and in
    four-row chunks (every chunk restarts the producer's cursors; every row pair straddles two chunks), and an odd
    split — dense (1080p), upscaling-ish small frames and sparse I420 staging — bit-exact against the oracle and
    identical from launch to launch."""
    import subprocess
    import sys
    env = dict(os.environ, TK_B200_PRE_DYN=dyn)
    r = subprocess.run([sys.executable, "-c", _DYN_CHILD], env=env, capture_output=True, text=True, timeout=600,
                       cwd=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    assert r.returncode == 0 and "DYN-OK" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]


def test_concurrent_launches_on_two_streams_take_different_chunk_counters():
    """Launches in flight at once each take their own chunk counter (g_pre_work[slot]): two streams, many launches,
    results identical to the serial ones."""
    n, w, h = 24, 1280, 720
    fa = torch.from_numpy(np.stack([synth.rgb_frame(w, h, 500 + i) for i in range(n)])).cuda()
    fb = torch.from_numpy(np.stack([synth.rgb_frame(w, h, 600 + i) for i in range(n)])).cuda()
    want_a = V.preprocess_batch(fa, 640, 640).clone()
    want_b = V.preprocess_batch(fb, 640, 640).clone()
    torch.cuda.synchronize()
    sa, sb = torch.cuda.Stream(), torch.cuda.Stream()
    oa = [torch.empty_like(want_a) for _ in range(6)]
    ob = [torch.empty_like(want_b) for _ in range(6)]
    for i in range(6):
        with torch.cuda.stream(sa):
            V.preprocess_batch(fa, 640, 640, out=oa[i])
        with torch.cuda.stream(sb):
            V.preprocess_batch(fb, 640, 640, out=ob[i])
    torch.cuda.synchronize()
    for i in range(6):
        assert torch.equal(oa[i], want_a) and torch.equal(ob[i], want_b)
