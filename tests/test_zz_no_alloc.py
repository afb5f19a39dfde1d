"""tk_b200_nms_classwise_static (csrc/nms.cu): the class-wise NMS with its scratch in a static device array, for callers
that cannot allocate.  Its output must equal tk_kernels_nms_classwise's byte for byte (integer / index work: the bar of
every NMS test here is bit-exact), on two calls in a row — the second call finds the first one's contents in the array.
Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest
import torch

import synth
from trackiellm_b200 import _capi as capi
from trackiellm_b200 import vision as V

pytestmark = pytest.mark.gpu


def _candidates(seeds, thr, cap, **kw):
    ts = np.stack([synth.logits_tensor(s, **kw) for s in seeds])
    c, n = V.decode_detections(torch.from_numpy(ts).cuda(), thr, max_candidates=cap)
    torch.cuda.synchronize()
    return c, n


def _both(c, n, iou, max_det, **kw):
    want_d, want_n = V.nms_classwise(c, n, iou, max_det, **kw)
    torch.cuda.synchronize()
    # outputs allocated (and poisoned) by the caller, before the call
    got_d = torch.full((c.shape[0], max_det, V.CAND_WORDS), -7, dtype=torch.int32, device=c.device)
    got_n = torch.full((c.shape[0],), -7, dtype=torch.int32, device=c.device)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    V.nms_classwise_static(c, n, iou, got_d, got_n, **kw)
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free_before           # nothing was allocated behind the caller's back
    return V.candidates_to_numpy(want_d, want_n), want_n.cpu().numpy(), V.candidates_to_numpy(got_d, got_n), got_n.cpu().numpy()


def test_static_scratch_nms_matches_the_allocating_entry_point_twice_in_a_row():
    assert V.nms_static_max_candidates() == 16384
    # call 1: crowded frames — more than 4096 candidates each, so the existing entry point takes its allocating path
    # (nms_global_kernel on stream-ordered scratch) and the static one fills its array with large work areas
    c1, n1 = _candidates([21, 22, 23], 0.2, 16384, objects=200, dups=30)
    assert int(n1.min()) > 4096
    w, wn, g, gn = _both(c1, n1, 0.5, 500)
    assert (wn == gn).all() and int(gn.min()) > 0
    for a, b in zip(w, g):
        assert a.tobytes() == b.tobytes()
    # call 2, straight after, other frames and another work-area size (512-candidate areas overlay what call 1 left
    # behind): ~480 candidates per frame, more frames than call 1, class-agnostic + coordinate restore on
    c2, n2 = _candidates(range(40, 52), 0.5, 512)
    assert 0 < int(n2.max()) <= 512
    w, wn, g, gn = _both(c2, n2, 0.45, 300, class_agnostic=True, restore=(1920, 1080, 640, 640))
    assert (wn == gn).all()
    for a, b in zip(w, g):
        assert a.tobytes() == b.tobytes()
    # call 3: call 1's input again — same answer as the first time
    w1, wn1, g1, gn1 = _both(c1, n1, 0.5, 500)
    assert (wn1 == gn1).all()
    for a, b in zip(w1, g1):
        assert a.tobytes() == b.tobytes()


def test_static_scratch_nms_rejects_what_it_has_no_room_for():
    c, n = _candidates([5], 0.5, 512)
    big = torch.zeros((1, 16385, V.CAND_WORDS), dtype=torch.int32, device="cuda")
    d = torch.zeros((1, 10, V.CAND_WORDS), dtype=torch.int32, device="cuda")
    nd = torch.zeros((1,), dtype=torch.int32, device="cuda")
    with pytest.raises(capi.TkError) as e:
        V.nms_classwise_static(big, n, 0.5, d, nd)
    assert e.value.code == capi.TK_ERROR_BUFFER_TOO_SMALL and "16384" in str(e.value)
    assert capi.lib().tk_b200_nms_classwise_static(None, None) == capi.TK_ERROR_INVALID_ARGUMENT
    # an empty batch is a no-op, as in tk_kernels_nms_classwise
    prm = capi.NmsParams(c.data_ptr(), n.data_ptr(), 0, 512, 0.5, 10, 0, d.data_ptr(), nd.data_ptr(), 0, 0, 0, 0)
    assert capi.lib().tk_b200_nms_classwise_static(C.byref(prm), None) == capi.TK_SUCCESS
