"""ctypes binding of libk3m_sm100a.so (the C ABI declared in include/k3m.h).

There is no CPU or library fallback: if the shared library is missing or a symbol fails to
resolve, importing the op table raises."""
import ctypes
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libk3m_sm100a.so")

LIB_F32_PATH = os.path.join(HERE, "libk3m_sm100a_f32.so")

_lib = None
_lib_f32 = None


class K3MError(RuntimeError):
    pass


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise K3MError(
            "k3m_b200: %s is missing — build it with `python -m k3m_b200.build` "
            "(nvcc, sm_100a). There is no fallback path." % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH)
    lib.k3m_last_error_string.restype = ctypes.c_char_p
    lib.k3m_get_version.restype = ctypes.c_int
    lib.k3m_gemm_last_variant.restype = ctypes.c_char_p
    _lib = lib
    return lib


def check(rc, what, lib=None):
    if rc != 0:
        msg = (lib or load()).k3m_last_error_string().decode("utf-8", "replace")
        raise K3MError("%s failed (%d): %s" % (what, rc, msg))


def load_f32():
    """The fp32 CHECK build (tests only): same entry points, fp32 activations / weights, CUDA-core GEMM / attention."""
    global _lib_f32
    if _lib_f32 is not None:
        return _lib_f32
    if not os.path.exists(LIB_F32_PATH):
        raise K3MError("k3m_b200: %s is missing — build it with `python -m k3m_b200.build`" % LIB_F32_PATH)
    lib = ctypes.CDLL(LIB_F32_PATH)
    lib.k3m_last_error_string.restype = ctypes.c_char_p
    lib.k3m_get_version.restype = ctypes.c_int
    lib.k3m_gemm_last_variant.restype = ctypes.c_char_p
    _lib_f32 = lib
    return lib
