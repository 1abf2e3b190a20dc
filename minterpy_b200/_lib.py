"""ctypes binding of libminterpy_b200.so (declared in include/minterpy_b200.h).

There is no CPU fallback: if the shared library is missing this module raises at import
of the first symbol, and every compute call needs a CUDA device.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libminterpy_b200.so")

c_i64 = ctypes.c_int64
c_int = ctypes.c_int
c_vp = ctypes.c_void_p

# name -> (restype, argtypes); must list every symbol include/minterpy_b200.h declares
SIGNATURES = {
    "mpx_version": (c_int, []),
    "mpx_last_error": (ctypes.c_char_p, []),
    "mpx_plan_create": (c_int, [c_int, c_vp, c_i64, c_int, c_vp, c_i64, ctypes.POINTER(c_vp)]),
    "mpx_plan_destroy": (None, [c_vp]),
    "mpx_plan_info": (c_int, [c_vp, c_int, ctypes.POINTER(c_i64)]),
    "mpx_eval": (c_int, [c_vp, c_vp, c_i64, c_i64, c_vp, c_int, c_i64, c_vp, c_i64, c_vp]),
    "mpx_eval_noalloc": (c_int, [c_vp, c_vp, c_i64, c_i64, c_vp, c_int, c_i64, c_vp, c_i64, c_vp]),
    "mpx_eval_host": (c_int, [c_vp, c_vp, c_i64, c_i64, c_vp, c_int, c_i64, c_vp, c_i64]),
    "mpx_monomials": (c_int, [c_vp, c_vp, c_i64, c_i64, c_vp, c_i64, c_int, c_vp]),
    "mpx_dds": (c_int, [c_vp, c_vp, c_int, c_i64, c_vp]),
    "mpx_dds_ex": (c_int, [c_vp, c_vp, c_int, c_i64, c_int, c_vp]),
    "mpx_dds_host": (c_int, [c_vp, c_vp, c_int, c_i64, c_vp, c_i64]),
    "mpx_apply_dense": (c_int, [c_int, c_vp, c_i64, c_i64, c_i64, c_vp, c_int, c_i64, c_vp,
                                c_i64, c_vp]),
    "mpx_apply_dense_ex": (c_int, [c_int, c_vp, c_i64, c_i64, c_i64, c_vp, c_int, c_i64, c_vp,
                                   c_i64, c_int, c_vp]),
    "mpx_eval_deriv": (c_int, [c_vp, c_vp, c_i64, c_i64, c_vp, c_int, c_i64, c_vp, c_vp, c_i64,
                               c_vp]),
    "mpx_index_is_sorted": (c_int, [c_vp, c_i64, c_int, ctypes.POINTER(c_int)]),
    "mpx_index_is_complete": (c_int, [c_vp, c_i64, c_int, ctypes.POINTER(c_int)]),
}
# not part of the public header: host-only view of the compiled program, used by tests
DEBUG_SIGNATURES = {
    "mpx_debug_program": (c_int, [c_vp, c_i64, c_int, c_vp, c_vp, ctypes.POINTER(c_i64),
                                  ctypes.POINTER(c_i64)]),
    "mpx_debug_tile_program": (c_int, [c_vp, c_i64, c_int, c_vp, c_vp, ctypes.POINTER(c_i64),
                                       ctypes.POINTER(c_i64)]),
    # test hooks: kernel choice that the library otherwise makes by itself (no environment
    # variable changes what the shipped library runs)
    "mpx_debug_set_tmma": (None, [c_int]),          # at plan creation: 0 auto, 1 force, 2 never
    "mpx_debug_set_dds_group": (None, [c_int]),     # dds through the fused slice kernel
    "mpx_debug_set_eval_stages": (None, [c_int]),   # ring depth of the tree kernel (0 = default)
    "mpx_debug_set_eval_uniform": (None, [c_int]),  # 2: record stream from the ring, never as a parameter
    "mpx_debug_set_eval_1d": (None, [c_int]),       # 2: m = 1 on the tree kernel, never Horner's rule
}

MPX_DENSE_LOWER = 1
MPX_DDS_MANY_ZEROS = 1
MPX_OK, MPX_ERR_ARG, MPX_ERR_INCOMPLETE, MPX_ERR_CUDA, MPX_ERR_UNSUPPORTED, MPX_ERR_NOMEM = range(6)
(INFO_N, INFO_M, INFO_COMPLETE, INFO_EVAL_FMA, INFO_DDS_UPDATES, INFO_LEAVES, INFO_DEVICE,
 INFO_SCRATCH_BYTES, INFO_EVAL_OTHER_FP64, INFO_CHUNKS, INFO_SMALLK_KERNEL,
 INFO_TMMA_FMA, INFO_MMA_MIN_K, INFO_UNIFORM_PROGRAM) = range(14)

_lib = None


class MinterpyB200Error(RuntimeError):
    """A CUDA / library failure (not an argument error)."""


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise MinterpyB200Error(
                f"{LIB_PATH} is missing: build it with `python -m minterpy_b200.build` "
                "(needs nvcc); there is no CPU fallback")
        handle = ctypes.CDLL(LIB_PATH)
        for table in (SIGNATURES, DEBUG_SIGNATURES):
            for name, (res, args) in table.items():
                fn = getattr(handle, name)
                fn.restype = res
                fn.argtypes = args
        _lib = handle
    return _lib


def last_error():
    msg = lib().mpx_last_error()
    return msg.decode() if msg else ""


def check(rc):
    """Map a status code to the exception the reference would raise."""
    if rc == MPX_OK:
        return
    msg = last_error()
    if rc in (MPX_ERR_ARG, MPX_ERR_INCOMPLETE):
        raise ValueError(msg)
    if rc == MPX_ERR_UNSUPPORTED:
        raise NotImplementedError(msg)
    if rc == MPX_ERR_NOMEM:
        raise MemoryError(msg)
    raise MinterpyB200Error(msg)
