"""b2r_backproject_static (include/b200recon.h): the allocation-free variant of b2r_backproject -- scratch in a file-scope
static __device__ array, count written to device memory, no context.  It has no counterpart in the reference,
so its check is the existing entry point: identical points / colours / count, on two
calls in a row (the second with a smaller image, so the scratch still holds the first call's offsets), with the
device's free memory unchanged across the calls.  Collected last on purpose (pytest -x)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _frame(rng, H, W, holes):
    d = rng.uniform(0.3, 2.5, (H, W)).astype(np.float32)
    d[rng.random((H, W)) < holes] = 0.0
    d[H // 3:H // 2, W // 4:W // 2] = 0.0
    rgb = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
    return d, rgb


def _both(o3d, d, rgb, K, ext):
    """-> (points, colours) of b2r_backproject and of b2r_backproject_static, and the free-memory change of the latter."""
    import torch
    from b200recon import _lib, runtime
    L = _lib.load()
    c = runtime.ctx(0)
    H, W = d.shape
    dev = c.tdev
    dd, dc = torch.from_numpy(d).to(dev), torch.from_numpy(rgb).to(dev)
    _, e16 = runtime.mat16(ext)
    Kc = _lib.Intrinsic(W, H, *K)
    # every buffer of both calls exists before the static call is issued
    p0 = torch.full((H * W, 3), np.nan, dtype=torch.float64, device=dev)
    c0 = torch.full((H * W, 3), np.nan, dtype=torch.float64, device=dev)
    p1, c1 = p0.clone(), c0.clone()
    n1 = torch.full((1,), -1, dtype=torch.int64, device=dev)
    n0 = C.c_int64(-1)
    _lib.check(L.b2r_backproject(c.handle, runtime.ptr(dd), runtime.ptr(dc), 3, C.byref(Kc), e16, runtime.ptr(p0),
                                 runtime.ptr(c0), C.byref(n0), c.stream()))
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info(dev)[0]
    _lib.check(L.b2r_backproject_static(dev.index or 0, runtime.ptr(dd), runtime.ptr(dc), 3, C.byref(Kc), e16, runtime.ptr(p1),
                                        runtime.ptr(c1), runtime.ptr(n1), c.stream()))
    torch.cuda.synchronize()
    free_after = torch.cuda.mem_get_info(dev)[0]
    n = int(n1.item())
    assert n == n0.value == int((d > 0).sum())
    return (p0[:n].cpu().numpy(), c0[:n].cpu().numpy()), (p1[:n].cpu().numpy(), c1[:n].cpu().numpy()), free_after - free_before


def test_backproject_static_equals_backproject_twice_in_a_row(o3d):
    from b200recon import _lib
    rng = np.random.default_rng(21)
    ext = np.array([[0.8, -0.6, 0.0, 0.1], [0.6, 0.8, 0.0, -0.2], [0.0, 0.0, 1.0, 0.3], [0.0, 0.0, 0.0, 1.0]])
    deltas = []
    # call 1: 640 x 480 (1200 scratch entries); call 2: 333 x 77, ragged against the 256-pixel blocks, 101 entries -- the
    # scratch behind them still holds call 1's offsets; call 3: the first image again
    for (H, W, holes) in ((480, 640, 0.2), (77, 333, 0.6), (480, 640, 0.05)):
        d, rgb = _frame(rng, H, W, holes)
        K = (600.0 * W / 640.0, 600.0 * W / 640.0, W / 2.0 - 0.5, H / 2.0 - 0.5)
        (p0, c0), (p1, c1), delta = _both(o3d, d, rgb, K, ext)
        assert len(p0) > 1000
        # the repo's back-projection tolerance is 1e-12 against the oracle; against the existing entry point the
        # arithmetic is the same kernel's, so the arrays are identical
        assert np.array_equal(p0, p1) and np.array_equal(c0, c1)
        assert np.abs(p0 - p1).max() <= 1e-12
        deltas.append(delta)
    # no allocation: the second and third static calls (the first may load the module's symbol table) leave the device's
    # free memory where it was
    assert deltas[1] == 0 and deltas[2] == 0, deltas


def test_backproject_static_all_invalid_and_limit(o3d):
    import torch
    from b200recon import _lib, runtime
    L = _lib.load()
    c = runtime.ctx(0)
    dev = c.tdev
    mx = C.c_int64(0)
    _lib.check(L.b2r_backproject_static_max_pixels(C.byref(mx)))
    assert mx.value == 1 << 22
    # no valid pixel: count 0, outputs untouched
    H, W = 9, 31
    dd = torch.zeros((H, W), dtype=torch.float32, device=dev)
    pts = torch.full((H * W, 3), 7.0, dtype=torch.float64, device=dev)
    n = torch.full((1,), -1, dtype=torch.int64, device=dev)
    _, e16 = runtime.mat16(np.eye(4))
    Kc = _lib.Intrinsic(W, H, 30.0, 30.0, 15.0, 4.0)
    _lib.check(L.b2r_backproject_static(dev.index or 0, runtime.ptr(dd), None, 0, C.byref(Kc), e16, runtime.ptr(pts), None,
                                        runtime.ptr(n), c.stream()))
    torch.cuda.synchronize()
    assert int(n.item()) == 0 and bool((pts == 7.0).all())
    # exactly the maximum is accepted by the size check, one row more is rejected with a clear error and nothing launched
    Kbig = _lib.Intrinsic(2048, 2049, 1000.0, 1000.0, 1024.0, 1024.0)
    rc = L.b2r_backproject_static(dev.index or 0, runtime.ptr(dd), None, 0, C.byref(Kbig), e16, runtime.ptr(pts), None,
                                  runtime.ptr(n), c.stream())
    assert rc == -7
    msg = L.b2r_last_error().decode()
    assert "exceed" in msg and "4194304" in msg, msg
    with pytest.raises(_lib.B2RError):
        _lib.check(rc)
    torch.cuda.synchronize()
    assert int(n.item()) == 0
