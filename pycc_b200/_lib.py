"""
ctypes binding of libpyccb200.so (C ABI: include/pyccb200.h).

The library is built in-tree by ``pycc_b200/csrc/build.sh`` (``__graft_entry__.build()``).
There is no CPU fallback: if the shared object is missing, or no B200 is visible, the engine
raises instead of computing anything.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libpyccb200.so")

F64, F32 = 0, 1
IMPL_TENSOR, IMPL_NAIVE = 0, 1
MAX_SEGMENTS = 4
ABI_VERSION = 1

# every symbol include/pyccb200.h declares (tests check the .so exports exactly these)
EXPORTS = (
    "pccb_abi_version", "pccb_device_count", "pccb_create", "pccb_destroy", "pccb_last_error",
    "pccb_set_stream", "pccb_set_workspace", "pccb_set_contract_impl", "pccb_launch_count",
    "pccb_synchronize", "pccb_contract", "pccb_pair_combine", "pccb_pair_symmetrize", "pccb_permute4",
    "pccb_update_amplitudes", "pccb_cc_energy", "pccb_dot", "pccb_axpby", "pccb_lincomb",
    "pccb_triples_energy", "pccb_t3_combine", "pccb_split_lo", "pccb_register_lo_plane", "pccb_clear_lo_planes",
    "pccb_unregister_lo_plane", "pccb_update_energy", "pccb_prepare_amplitudes",
    "pccb_cc_energy_static", "pccb_cc_energy_static_max_partials", "pccb_dot_many", "pccb_t3_density_combine",
)


class CIdx(C.Structure):
    _fields_ = [("div", C.c_int64), ("s_hi", C.c_int64), ("s_lo", C.c_int64)]


class Segment(C.Structure):
    _fields_ = [("A_dev", C.c_void_p), ("B_dev", C.c_void_p),
                ("am", CIdx), ("ak", CIdx), ("bn", CIdx), ("bk", CIdx),
                ("K", C.c_int64), ("a_batch", C.c_int64), ("b_batch", C.c_int64),
                ("scale", C.c_double)]


class Contraction(C.Structure):
    _fields_ = [("M", C.c_int64), ("N", C.c_int64), ("batch", C.c_int64),
                ("nseg", C.c_int32), ("dtype", C.c_int32),
                ("seg", Segment * MAX_SEGMENTS),
                ("C_dev", C.c_void_p), ("cm", CIdx), ("cn", CIdx), ("c_batch", C.c_int64),
                ("D_dev", C.c_void_p), ("dm", CIdx), ("dn", CIdx), ("d_batch", C.c_int64),
                ("alpha", C.c_double), ("beta", C.c_double), ("gamma", C.c_double)]


class PrepareOutputs(C.Structure):
    _fields_ = [(k, C.c_void_p) for k in ("tau", "tauh", "tauW", "t2s", "t11", "t2sP", "t2P", "t2Q", "tauWPn", "t2Rh")]


class EngineError(RuntimeError):
    pass


_lib = None


def load():
    """dlopen libpyccb200.so and declare the prototypes.  Raises EngineError when it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise EngineError("libpyccb200.so not found at %s: run pycc_b200/csrc/build.sh "
                          "(or __graft_entry__.build()); there is no CPU fallback" % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    vp, i32, i64, dbl = C.c_void_p, C.c_int, C.c_int64, C.c_double
    pd = C.POINTER(C.c_double)
    lib.pccb_abi_version.restype = i32
    lib.pccb_device_count.restype = i32
    lib.pccb_create.argtypes = [i32, C.POINTER(vp)]
    lib.pccb_destroy.argtypes = [vp]
    lib.pccb_last_error.argtypes = [vp]
    lib.pccb_last_error.restype = C.c_char_p
    lib.pccb_set_stream.argtypes = [vp, vp]
    lib.pccb_set_workspace.argtypes = [vp, vp, C.c_size_t]
    lib.pccb_set_contract_impl.argtypes = [vp, i32]
    lib.pccb_launch_count.argtypes = [vp, C.POINTER(i64)]
    lib.pccb_synchronize.argtypes = [vp]
    lib.pccb_contract.argtypes = [vp, C.POINTER(Contraction)]
    lib.pccb_pair_combine.argtypes = [vp, i32, vp, vp, dbl, dbl, vp, vp, dbl, i64, i64]
    lib.pccb_pair_symmetrize.argtypes = [vp, i32, vp, i64, i64]
    lib.pccb_permute4.argtypes = [vp, i32, vp, vp, C.POINTER(i64), C.POINTER(i64), dbl, dbl]
    lib.pccb_update_amplitudes.argtypes = [vp, i32, vp, vp, vp, vp, vp, i64, i64, i32, pd]
    lib.pccb_update_energy.argtypes = [vp, i32, vp, vp, vp, vp, vp, vp, i64, vp, i64, i64, i32, i64, i64, i32, pd]
    lib.pccb_prepare_amplitudes.argtypes = [vp, i32, vp, vp, i64, i64, i64, i64, C.POINTER(PrepareOutputs)]
    lib.pccb_cc_energy_static.argtypes = [vp, i32, vp, i64, vp, vp, vp, i64, i64, vp]
    lib.pccb_cc_energy.argtypes = [vp, i32, vp, i64, vp, vp, vp, i64, i64, pd]
    lib.pccb_dot.argtypes = [vp, i32, vp, vp, i64, pd]
    lib.pccb_t3_density_combine.argtypes = [vp, i32, vp, vp, i32, i32, i32, i32, vp, i64, i64, C.POINTER(vp)]
    lib.pccb_dot_many.argtypes = [vp, i32, vp, C.POINTER(vp), i32, i64, pd]
    lib.pccb_axpby.argtypes = [vp, i32, vp, dbl, vp, dbl, vp, i64]
    lib.pccb_lincomb.argtypes = [vp, i32, vp, C.POINTER(vp), pd, i32, i64]
    lib.pccb_triples_energy.argtypes = [vp, i32, vp, vp, i32, vp, vp, vp, vp, i64, vp, i64, i64, vp]
    lib.pccb_split_lo.argtypes = [vp, vp, vp, i64]
    lib.pccb_register_lo_plane.argtypes = [vp, vp, vp, i64]
    lib.pccb_clear_lo_planes.argtypes = [vp]
    lib.pccb_unregister_lo_plane.argtypes = [vp, vp]
    lib.pccb_t3_combine.argtypes = [vp, i32, vp, i32, i32, i32, i32, vp, i64, i64, vp, vp]
    for name in EXPORTS:
        fn = getattr(lib, name)
        if name == "pccb_cc_energy_static_max_partials":
            fn.restype = i64
        elif name not in ("pccb_last_error",):
            fn.restype = i32
    if lib.pccb_abi_version() != ABI_VERSION:
        raise EngineError("libpyccb200.so ABI version mismatch")
    _lib = lib
    return lib
