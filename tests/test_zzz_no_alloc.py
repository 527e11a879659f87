"""fod_set_loss_static — the allocation-free variant of fod_set_loss (include/fod_b200.h): its (L,B,8) partial-sum
workspace is a file-scope device array inside the library instead of caller memory.  Called here straight through the
C ABI (ctypes), checked against the existing entry point on two calls in a row with different inputs (the array keeps
the previous call's contents), plus the size limit.  Tolerance: 1e-4, as tests/test_criterion_gpu.py uses for the loss
path.  Collected last on purpose (file name): the round-end tier runs `pytest -x`."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _inputs(seed, levels, bsz, m=128, c=8, nt_cap=40):
    g = torch.Generator().manual_seed(seed)
    logits = (torch.randn(levels, bsz, m, c, generator=g) * 2 - 2).to(DEV)
    boxes = torch.cat((torch.rand(levels, bsz, m, 2, generator=g), torch.rand(levels, bsz, m, 2, generator=g) * 0.5 + 0.01), -1).to(DEV)
    cnt = torch.randint(0, nt_cap, (bsz,), generator=g, dtype=torch.int32)
    tb = torch.cat((torch.rand(bsz, nt_cap, 2, generator=g), torch.rand(bsz, nt_cap, 2, generator=g) * 0.4 + 0.01), -1)
    tl = torch.randint(0, c, (bsz, nt_cap), generator=g, dtype=torch.int32)
    for b in range(bsz):
        tb[b, int(cnt[b]):] = 0
        tl[b, int(cnt[b]):] = 0
    return logits, boxes, tb.to(DEV), tl.to(DEV), cnt.to(DEV)


def _run(entry, logits, boxes, match_tgt, tb, tl, cnt, nb, partial):
    """One call of `entry` (a ctypes function of the library) on the current stream; outputs are the caller's."""
    from fod_b200 import lib
    levels, bsz, m, c = logits.shape
    out = torch.zeros(8 + 3 * levels, dtype=torch.float32, device=DEV)
    dl, db = torch.empty_like(logits), torch.empty_like(boxes)
    a = lib.SetLossArgs()
    p = lambda t: None if t is None else C.c_void_p(t.data_ptr())  # noqa: E731
    a.logits, a.boxes, a.match_tgt = p(logits), p(boxes), p(match_tgt)
    a.tgt_boxes, a.tgt_labels, a.tgt_count, a.num_boxes = p(tb), p(tl), p(cnt), p(nb)
    a.dlogits, a.dboxes, a.partial, a.out = p(dl), p(db), p(partial), p(out)
    a.L, a.B, a.M, a.C, a.NT, a.main_level = levels, bsz, m, c, tb.shape[1], levels - 1
    a.w_ce, a.w_bbox, a.w_giou, a.alpha = 2.0, 5.0, 2.0, 0.25
    rc = entry(C.byref(a), C.c_void_p(torch.cuda.current_stream().cuda_stream))
    return rc, out, dl, db


def _close(a, b, what):
    err = float((a - b).abs().max() / max(1.0, float(b.abs().max())))
    assert err < 1e-4, (what, err)


def test_set_loss_static_matches_set_loss_on_two_calls_in_a_row():
    from fod_b200 import lib, ops
    L = lib.load()
    # the second call is smaller than the first: stale partial sums of call 1 lie beyond (and inside) its region
    for seed, levels, bsz in ((11, 6, 5), (12, 3, 2)):
        logits, boxes, tb, tl, cnt = _inputs(seed, levels, bsz)
        match_tgt = ops.match(logits, boxes, tb, tl, cnt)["match_tgt"]
        nb = torch.tensor([max(1.0, float(cnt.sum()))], device=DEV)
        partial = torch.empty((levels, bsz, 8), dtype=torch.float32, device=DEV)
        rc, out_ref, dl_ref, db_ref = _run(L.fod_set_loss, logits, boxes, match_tgt, tb, tl, cnt, nb, partial)
        assert rc == 0, L.fod_last_error()
        rc, out, dl, db = _run(L.fod_set_loss_static, logits, boxes, match_tgt, tb, tl, cnt, nb, None)
        assert rc == 0, L.fod_last_error()
        torch.cuda.synchronize()
        assert bool(torch.isfinite(out).all()) and float(out[0]) > 0
        _close(out, out_ref, "out")
        _close(dl, dl_ref, "dlogits")
        _close(db, db_ref, "dboxes")


def test_set_loss_static_rejects_sizes_beyond_its_scratch():
    from fod_b200 import lib
    L = lib.load()
    logits, boxes, tb, tl, cnt = _inputs(13, 2, 2)
    nb = torch.ones(1, device=DEV)
    mt = torch.full((2, 2, 128), -1, dtype=torch.int32, device=DEV)
    a = lib.SetLossArgs()
    p = lambda t: C.c_void_p(t.data_ptr())  # noqa: E731
    out = torch.zeros(8 + 3 * 16, device=DEV)
    a.logits, a.boxes, a.match_tgt = p(logits), p(boxes), p(mt)
    a.tgt_boxes, a.tgt_labels, a.tgt_count, a.num_boxes, a.out = p(tb), p(tl), p(cnt), p(nb), p(out)
    a.L, a.B, a.M, a.C, a.NT, a.main_level = 16, 257, 128, 8, 40, 15   # 4112 > 4096: rejected before any launch
    a.w_ce, a.w_bbox, a.w_giou, a.alpha = 2.0, 5.0, 2.0, 0.25
    rc = L.fod_set_loss_static(C.byref(a), C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 1  # FOD_E_ARG
    msg = L.fod_last_error()
    msg = msg.decode() if isinstance(msg, bytes) else str(msg)
    assert "FOD_SET_LOSS_STATIC_MAX_LB" in msg, msg
    torch.cuda.synchronize()
