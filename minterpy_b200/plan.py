"""`Plan`: the device-resident description of one polynomial space (index set + grid).

Wraps ``mpx_plan_*`` / ``mpx_eval*`` / ``mpx_dds*`` / ``mpx_monomials`` of the C ABI.  NumPy
arrays go through the ``*_host`` entry points (the library does the transfers); CUDA torch
tensors are passed by ``data_ptr()`` without a copy and results come back as torch tensors.
"""
import ctypes
import threading
import zlib

import numpy as np

from . import _lib

_cache = {}
_cache_lock = threading.Lock()
_CACHE_MAX = 32


def _is_torch(a):
    return type(a).__module__.split(".")[0] == "torch"


def _current_device():
    import torch
    if not torch.cuda.is_available():
        raise _lib.MinterpyB200Error("no CUDA device available; minterpy_b200 has no CPU fallback")
    return torch.cuda.current_device()


def _host_result(shape):
    """Result array for the NumPy path, backed by page-locked memory from torch's caching host
    allocator (a device-to-host copy into pageable memory is several times slower, and the
    allocator recycles the block once the array is garbage collected)."""
    n = int(np.prod(shape))
    if n * 8 < (1 << 20):
        return np.empty(shape, dtype=np.float64)
    import torch
    return torch.empty(shape, dtype=torch.float64, pin_memory=True).numpy()


class Plan:
    def __init__(self, exponents, generating_points, device=None):
        e = np.ascontiguousarray(exponents, dtype=np.int64)
        gp = np.ascontiguousarray(generating_points, dtype=np.float64)
        if e.ndim != 2 or gp.ndim != 2:
            raise ValueError("exponents and generating_points must be 2-D")
        if gp.shape[1] != e.shape[1]:
            raise ValueError(
                f"generating_points has {gp.shape[1]} columns but the exponents have {e.shape[1]}")
        if device is None:
            device = _current_device()
        self.device = int(device)
        self.N, self.m = e.shape
        handle = ctypes.c_void_p()
        rc = _lib.lib().mpx_plan_create(self.device, e.ctypes.data, self.N, self.m,
                                        gp.ctypes.data, gp.shape[0], ctypes.byref(handle))
        _lib.check(rc)
        self._h = handle
        # private copies: the cache compares them with the caller's arrays on every hit
        self.exponents = e.copy()
        self.generating_points = gp.copy()

    def same_space(self, e, gp):
        return (self.exponents.shape == e.shape and self.generating_points.shape == gp.shape and
                np.array_equal(self.exponents, e) and np.array_equal(self.generating_points, gp))

    def __del__(self):
        h = getattr(self, "_h", None)
        if h:
            try:
                _lib.lib().mpx_plan_destroy(h)
            except Exception:
                pass
            self._h = None

    def info(self, what):
        v = ctypes.c_int64()
        _lib.check(_lib.lib().mpx_plan_info(self._h, what, ctypes.byref(v)))
        return v.value

    @property
    def is_complete(self):
        return bool(self.info(_lib.INFO_COMPLETE))

    # ---- evaluation ---------------------------------------------------------------------
    def eval(self, xx, coeffs):
        """xx (P, m), coeffs (N, k) -> (P, k).  NumPy in -> NumPy out, torch in -> torch out."""
        if _is_torch(xx) or _is_torch(coeffs):
            return self._eval_torch(xx, coeffs)
        xx = np.ascontiguousarray(xx, dtype=np.float64)
        coeffs = np.ascontiguousarray(coeffs, dtype=np.float64)
        P, k = xx.shape[0], coeffs.shape[1]
        out = _host_result((P, k))
        if P:
            rc = _lib.lib().mpx_eval_host(self._h, xx.ctypes.data, P, self.m, coeffs.ctypes.data,
                                          k, k, out.ctypes.data, k)
            _lib.check(rc)
        return out

    def eval_into(self, xx, coeffs, out):
        """NumPy path with a caller-provided result: xx (P, m), coeffs (N, k), out (P, k), all
        float64 with unit stride along rows (slices of larger arrays are fine)."""
        for name, a in (("xx", xx), ("coeffs", coeffs), ("out", out)):
            if a.dtype != np.float64 or a.ndim != 2 or a.strides[1] != 8:
                raise ValueError(f"{name} must be a 2-D float64 array with contiguous rows")
        P, k = xx.shape[0], coeffs.shape[1]
        if out.shape != (P, k) or xx.shape[1] != self.m or coeffs.shape[0] != self.N:
            raise ValueError("shape mismatch between points, coefficients and result")
        if P:
            rc = _lib.lib().mpx_eval_host(self._h, xx.ctypes.data, P, xx.strides[0] // 8,
                                          coeffs.ctypes.data, k, coeffs.strides[0] // 8,
                                          out.ctypes.data, out.strides[0] // 8)
            _lib.check(rc)
        return out

    def _eval_torch(self, xx, coeffs, out=None):
        import torch
        dev = torch.device("cuda", self.device)
        xx = torch.as_tensor(xx, dtype=torch.float64, device=dev)
        coeffs = torch.as_tensor(coeffs, dtype=torch.float64, device=dev)
        if xx.stride(-1) != 1:
            xx = xx.contiguous()
        if coeffs.stride(-1) != 1:
            coeffs = coeffs.contiguous()
        P, k = xx.shape[0], coeffs.shape[1]
        if out is None:
            out = torch.empty((P, k), dtype=torch.float64, device=dev)
        if P:
            stream = torch.cuda.current_stream(dev).cuda_stream
            rc = _lib.lib().mpx_eval(self._h, xx.data_ptr(), P, xx.stride(0), coeffs.data_ptr(), k,
                                     coeffs.stride(0), out.data_ptr(), out.stride(0), stream)
            _lib.check(rc)
        return out

    def eval_deriv(self, xx, coeffs, orders):
        """Partial derivative of the given orders (m ints): xx (P, m), coeffs (N, k) -> (P, k).
        The device work is ``mpx_eval_deriv``; NumPy inputs are staged through torch."""
        import torch
        dev = torch.device("cuda", self.device)
        as_numpy = not (_is_torch(xx) or _is_torch(coeffs))
        orders = np.ascontiguousarray(orders, dtype=np.int32).reshape(-1)
        if orders.shape[0] != self.m:
            raise ValueError(f"{self.m} derivative orders expected, {orders.shape[0]} given")
        x = torch.as_tensor(np.ascontiguousarray(xx) if not _is_torch(xx) else xx,
                            dtype=torch.float64, device=dev)
        c = torch.as_tensor(np.ascontiguousarray(coeffs) if not _is_torch(coeffs) else coeffs,
                            dtype=torch.float64, device=dev)
        if x.stride(-1) != 1:
            x = x.contiguous()
        if c.stride(-1) != 1:
            c = c.contiguous()
        P, k = x.shape[0], c.shape[1]
        out = torch.empty((P, k), dtype=torch.float64, device=dev)
        if P:
            stream = torch.cuda.current_stream(dev).cuda_stream
            rc = _lib.lib().mpx_eval_deriv(self._h, x.data_ptr(), P, x.stride(0), c.data_ptr(), k,
                                           c.stride(0), orders.ctypes.data, out.data_ptr(),
                                           out.stride(0), stream)
            _lib.check(rc)
        return out.cpu().numpy() if as_numpy else out

    def monomials(self, xx, triangular=False):
        """(P, N) Newton monomial matrix."""
        import torch
        dev = torch.device("cuda", self.device)
        as_numpy = not _is_torch(xx)
        x = torch.as_tensor(np.ascontiguousarray(xx) if as_numpy else xx, dtype=torch.float64,
                            device=dev).contiguous()
        P = x.shape[0]
        out = torch.empty((P, self.N), dtype=torch.float64, device=dev)
        if P:
            stream = torch.cuda.current_stream(dev).cuda_stream
            rc = _lib.lib().mpx_monomials(self._h, x.data_ptr(), P, x.stride(0), out.data_ptr(),
                                          out.stride(0), int(bool(triangular)), stream)
            _lib.check(rc)
        return out.cpu().numpy() if as_numpy else out

    # ---- divided differences ----------------------------------------------------------------
    def dds(self, values, many_zeros=False, inplace=False):
        """values (N, k) -> Newton coefficients (N, k); the input is never modified, unless it
        is a contiguous float64 tensor on the plan's device and ``inplace`` is set.
        ``many_zeros``: the ``MPX_DDS_MANY_ZEROS`` hint of ``mpx_dds_ex`` (columns of the identity
        and the like; same bits out, another instantiation of the kernel)."""
        if _is_torch(values) or many_zeros:
            import torch
            dev = torch.device("cuda", self.device)
            as_numpy = not _is_torch(values)
            if as_numpy:        # a fresh device copy
                v = torch.as_tensor(np.ascontiguousarray(values, dtype=np.float64), device=dev)
            else:
                v = torch.as_tensor(values, dtype=torch.float64, device=dev)
                if not (inplace and v.is_contiguous() and v.data_ptr() == values.data_ptr()):
                    v = v.clone(memory_format=torch.contiguous_format)
            stream = torch.cuda.current_stream(dev).cuda_stream
            rc = _lib.lib().mpx_dds_ex(self._h, v.data_ptr(), v.shape[1], v.stride(0),
                                       _lib.MPX_DDS_MANY_ZEROS if many_zeros else 0, stream)
            _lib.check(rc)
            return v.cpu().numpy() if as_numpy else v
        values = np.ascontiguousarray(values, dtype=np.float64)
        out = _host_result(values.shape)      # page-locked for large results: faster download
        k = values.shape[1]
        rc = _lib.lib().mpx_dds_host(self._h, values.ctypes.data, k, k, out.ctypes.data, k)
        _lib.check(rc)
        return out


def get_plan(exponents, generating_points, device=None):
    """Plan for (exponents, generating_points) on `device`, cached by content."""
    e = np.ascontiguousarray(exponents, dtype=np.int64)
    gp = np.ascontiguousarray(generating_points, dtype=np.float64)
    if device is None:
        device = _current_device()
    # The key is only a hash (two 32-bit checksums per array, computed without copying); a hit
    # is confirmed by comparing the arrays with the plan's own copies, so a checksum collision
    # costs a rebuild, never a wrong plan.
    key = (int(device), e.shape, gp.shape, zlib.crc32(e), zlib.adler32(e), zlib.crc32(gp),
           zlib.adler32(gp))
    with _cache_lock:
        plan = _cache.get(key)
        if plan is not None and not plan.same_space(e, gp):
            plan = None
        if plan is None:
            plan = Plan(e, gp, device)
            if key not in _cache and len(_cache) >= _CACHE_MAX:
                _cache.pop(next(iter(_cache)))
            _cache[key] = plan
    return plan


def clear_plan_cache():
    with _cache_lock:
        _cache.clear()
