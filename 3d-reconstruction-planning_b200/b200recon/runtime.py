"""Device runtime: one library context per (process, device); PyTorch only owns device memory and
streams (tensors are passed to the C ABI as raw `data_ptr()`s)."""
import ctypes as C
import os

import numpy as np

from . import _lib

_ctxs = {}


def torch():
    import torch as _t
    return _t


def default_device():
    """One process per GPU: LOCAL_RANK selects the device (torchrun), else cuda:0."""
    return int(os.environ.get("LOCAL_RANK", "0"))


class Context:
    def __init__(self, device):
        t = torch()
        if not t.cuda.is_available():
            raise _lib.B2RError("b200recon needs a CUDA device (B200, sm_100a); none is visible and there is no "
                                "CPU fallback")
        self.device = int(device)
        self.tdev = t.device("cuda", self.device)
        t.cuda.set_device(self.device)
        t.zeros(1, device=self.tdev)  # make sure the primary context exists
        h = C.c_void_p()
        _lib.check(_lib.load().b2r_ctx_create(self.device, C.byref(h)))
        self.handle = h
        self.updates = t.zeros(1, dtype=t.int64, device=self.tdev)

    def stream(self):
        return C.c_void_p(torch().cuda.current_stream(self.tdev).cuda_stream)

    def launch_count(self):
        n = C.c_int64(0)
        _lib.check(_lib.load().b2r_ctx_launch_count(self.handle, C.byref(n)))
        return n.value

    def kernel_timing(self, enable):
        _lib.check(_lib.load().b2r_ctx_kernel_timing(self.handle, int(bool(enable))))

    def kernel_time(self):
        """(total ms, launches) of the tsdf_integrate kernel since kernel_timing(True) (CUDA events)."""
        ms, n = C.c_double(0), C.c_int64(0)
        _lib.check(_lib.load().b2r_ctx_kernel_time(self.handle, C.byref(ms), C.byref(n)))
        return ms.value, n.value

    def trace(self, max_calls):
        """Start (max_calls > 0) / stop (0) the device-side timeline of integrate calls (b2r_ctx_trace)."""
        _lib.check(_lib.load().b2r_ctx_trace(self.handle, int(max_calls)))

    def trace_read(self, cap=4096):
        """([n, 4, 2] uint64 ns: {start, end} of frame_prepare / integrate_plan / integrate_items / integrate_cull per
        recorded call, [n] live tiles, [n] work items)."""
        import numpy as np
        buf = (C.c_uint64 * (cap * 16))()
        n = C.c_int(0)
        _lib.check(_lib.load().b2r_ctx_trace_read(self.handle, buf, cap, C.byref(n)))
        a = np.frombuffer(buf, dtype=np.uint64)[: n.value * 16].reshape(n.value, 16).copy()
        self.trace_raw = a      # col 9: first streaming warp out of work (ns); col 10: sum over warps of (dry time - start)
        return a[:, :8].reshape(n.value, 4, 2), (a[:, 8] >> np.uint64(32)).astype(np.int64), (a[:, 8] & np.uint64(0xffffffff)).astype(np.int64)

    def join(self):
        """Order the current torch stream after every integration queued so far (their streaming kernels run on
        a stream owned by the library context).  Needed only before touching a volume's tensors directly."""
        _lib.check(_lib.load().b2r_ctx_join(self.handle, self.stream()))

    def updates_mirror(self):
        """Pinned-host copy of the update counter of the newest completed integrate call (see b200recon.h)."""
        n = C.c_uint64(0)
        _lib.check(_lib.load().b2r_ctx_updates_mirror(self.handle, C.byref(n)))
        return n.value

    def sync(self):
        _lib.check(_lib.load().b2r_ctx_sync(self.handle, self.stream()))


def ctx(device=None):
    d = default_device() if device is None else int(device)
    c = _ctxs.get(d)
    if c is None:
        c = Context(d)
        _ctxs[d] = c
    return c


def new_context(device=None):
    """A further library context on `device` (own scratch, own side/copy streams, own extraction state): one per
    slab when several slabs of a cube live in ONE process (parallel.LocalSlabs)."""
    return Context(default_device() if device is None else int(device))


def ptr(tensor):
    return C.c_void_p(0 if tensor is None else tensor.data_ptr())


def mat16(m):
    a = np.ascontiguousarray(np.asarray(m, dtype=np.float64).reshape(4, 4))
    return a, a.ctypes.data_as(_lib.c_dp)


def to_device(array, dtype=None, device=None):
    """numpy (or tensor) -> contiguous CUDA tensor on the context's device."""
    t = torch()
    c = ctx(device)
    if isinstance(array, t.Tensor):
        x = array
        if dtype is not None:
            x = x.to(dtype)
        return x.to(c.tdev).contiguous()
    a = np.ascontiguousarray(array)
    x = t.from_numpy(a)
    if dtype is not None:
        x = x.to(dtype)
    return x.to(c.tdev)
