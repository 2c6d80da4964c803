"""vslamb200_interest_points_noalloc (no counterpart in the reference):
the allocation-free variant of the interest-list construction, through the C ABI, against the lists the existing entry
point builds (vslamb200_evaluate's residual plane marks exactly the retained interest points).  Index work: the lists
must be identical.  Two calls in a row share the library's static scratch array; the second must not see the first's
contents.  In a file of its own that pytest collects last."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _call(A, torch, planes, depth_key, W, H, min_gradient):
    gx = torch.from_numpy(planes["dIx"]).cuda()
    gy = torch.from_numpy(planes["dIy"]).cuda()
    d = torch.from_numpy(planes[depth_key]).cuda()
    pts = torch.full((W * H,), -1, dtype=torch.int32, device="cuda")  # uint32 indices, viewed as such on the host
    cnt = torch.full((1,), -1, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    rc = A.lib().vslamb200_interest_points_noalloc(C.c_void_p(gx.data_ptr()), C.c_void_p(gy.data_ptr()), C.c_void_p(d.data_ptr()),
                                                   W, H, float(min_gradient), C.c_void_p(pts.data_ptr()),
                                                   C.c_void_p(cnt.data_ptr()), None)
    A.check(rc)
    torch.cuda.synchronize()
    n = int(cnt.cpu()[0])
    return n, pts.cpu().numpy().view(np.uint32)[:max(n, 0)].astype(np.int64)


def test_interest_points_noalloc_matches_the_allocating_path_twice(capi, oracle, synth):
    torch = pytest.importorskip("torch")
    A, O = capi, oracle
    W, H, L = 640, 480, 3
    p = synth.make_pair(W, H, 4242, holes_permille=60)
    ctx = A.Context(A.make_config(min_gradient=30.0))
    try:
        ctx.upload(np.array([1, 2]), np.stack([p["gray0"], p["gray1"]]), np.stack([p["depth0"], p["depth1"]]), p["cam"], L)
        # two calls in a row on different levels and masks: level 0 with the single-frame mask, level 1 with the
        # multi-frame one; then level 0 again (the scratch array now holds level 1's records)
        for level, multi in ((0, False), (1, True), (0, False)):
            ref = ctx.evaluate(1, 2, level, O.IDENTITY, multi_mask=multi)
            want = np.flatnonzero(ref["rcode"].ravel() != A.R_NOT_INTEREST)
            planes = ctx.download(1, level)
            Wl, Hl = planes["dIx"].shape[1], planes["dIx"].shape[0]
            n, got = _call(A, torch, planes, "Z" if multi else "Zsf", Wl, Hl, 30.0)
            assert n == ref["n"] == len(want), (level, n, ref["n"], len(want))
            assert np.array_equal(got, want), level
    finally:
        ctx.close()


def test_interest_points_noalloc_rejects_sizes_beyond_its_static_scratch(capi):
    torch = pytest.importorskip("torch")
    A = capi
    t = torch.zeros(64, dtype=torch.int32, device="cuda")
    ptr = C.c_void_p(t.data_ptr())
    rc = A.lib().vslamb200_interest_points_noalloc(ptr, ptr, ptr, 4097, 4096, 30.0, ptr, ptr, None)
    assert rc == -1  # VSLAMB200_EINVAL
    assert b"VSLAMB200_NOALLOC_MAX_PIXELS" in A.lib().vslamb200_last_error()
    assert A.lib().vslamb200_interest_points_noalloc(None, ptr, ptr, 8, 8, 30.0, ptr, ptr, None) == -1
