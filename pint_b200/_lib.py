"""ctypes binding of libpint_cuda.so (include/pint_cuda.h).

The library is the product; there is no fallback.  If it is missing this module
raises at import of the first symbol -- build it with `make` (or
`python -c "import __graft_entry__ as g; g.build()"`).
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "lib", "libpint_cuda.so")
HEADER_PATH = os.path.join(os.path.dirname(HERE), "include", "pint_cuda.h")

MAX_FIELDS = 64


class PintLayout(C.Structure):
    _fields_ = [("elem_bytes", C.c_uint32), ("nfields", C.c_uint32), ("bits", C.c_uint8 * MAX_FIELDS)]


class PintCudaError(RuntimeError):
    def __init__(self, rc, what):
        self.rc = rc
        super().__init__(f"{what}: rc={rc} ({error_string(rc)})")


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: the CUDA library is the only implementation (no CPU fallback). "
                "Build it with `make` at the repo root.")
        L = C.CDLL(LIB_PATH)
        vp, sz, u32, u64, i = C.c_void_p, C.c_size_t, C.c_uint32, C.c_uint64, C.c_int
        LP = C.POINTER(PintLayout)
        sigs = {
            "pint_cuda_abi_version": (i, []),
            "pint_cuda_error_string": (C.c_char_p, [i]),
            "pint_cuda_launch_count": (C.c_ulonglong, []),
            "pint_cuda_set_option": (i, [C.c_char_p, C.c_long]),
            "pint_cuda_get_option": (C.c_long, [C.c_char_p]),
            "pint_cuda_layout_validate": (i, [LP]),
            "pint_cuda_storage_bytes": (u32, [u32]),
            "pint_cuda_layout_masks": (i, [LP, C.POINTER(u64), C.POINTER(u64), C.POINTER(u64), C.POINTER(u32)]),
            "pint_cuda_layout_is_specialized": (i, [LP]),
            "pint_cuda_jit_compile": (i, [LP]),
            "pint_cuda_jit_prebuild": (i, [LP, i, i]),
            "pint_cuda_jit_last_error": (C.c_char_p, []),
            "pint_cuda_slice_layout": (i, [LP, u32, u32, LP]),
            "pint_cuda_binary": (i, [i, LP, vp, vp, vp, sz, vp]),
            "pint_cuda_fused_unpack_arith_repack": (i, [i, LP, vp, vp, vp, sz, vp]),
            "pint_cuda_binary_multi": (i, [i, C.POINTER(i), LP, vp, vp, C.POINTER(vp), sz, vp]),
            "pint_cuda_shift": (i, [i, LP, u64, vp, vp, sz, vp]),
            "pint_cuda_slice": (i, [LP, u32, u32, vp, vp, sz, vp]),
            "pint_cuda_unpack": (i, [LP, i, vp, C.POINTER(vp), i, sz, vp]),
            "pint_cuda_repack": (i, [LP, C.POINTER(vp), i, vp, sz, vp]),
            "pint_cuda_dense_bytes": (sz, [LP, sz]),
            "pint_cuda_dense_binary": (i, [i, LP, vp, vp, vp, sz, vp]),
            "pint_cuda_dense_binary_multi": (i, [i, C.POINTER(i), LP, vp, vp, C.POINTER(vp), sz, vp]),
            "pint_cuda_dense_shift": (i, [i, LP, u64, vp, vp, sz, vp]),
            "pint_cuda_dense_pack": (i, [LP, vp, vp, sz, vp]),
            "pint_cuda_dense_unpack": (i, [LP, vp, vp, sz, vp]),
            "pint_cuda_dense_to_planes": (i, [LP, i, vp, C.POINTER(vp), i, sz, vp]),
            "pint_cuda_dense_from_planes": (i, [LP, C.POINTER(vp), i, vp, sz, vp]),
            "pint_cuda_dense_plan_shards": (i, [sz, i, C.POINTER(sz), C.POINTER(sz)]),
            "pint_cuda_sharded_dense_binary": (i, [i, LP, i, C.POINTER(i), C.POINTER(vp), C.POINTER(vp), C.POINTER(vp),
                                                  C.POINTER(sz), C.POINTER(vp)]),
            "pint_cuda_pairs_reduce": (i, [i, LP, vp, sz, vp, vp]),
            "pint_cuda_pairs_reduce_static": (i, [i, LP, vp, sz, vp, vp]),
            "pint_cuda_equal": (i, [LP, vp, vp, sz, vp, vp, vp]),
            "pint_cuda_binary_pairs": (i, [i, LP, vp, vp, sz, vp]),
            "pint_cuda_binary_scalar": (i, [i, LP, vp, u64, i, vp, sz, vp]),
            "pint_cuda_fill_splitmix": (i, [vp, u32, sz, u64, u64, vp]),
            "pint_cuda_checksum": (i, [vp, u32, sz, vp, vp]),
            "pint_cuda_plan_shards": (i, [sz, u32, i, C.POINTER(sz), C.POINTER(sz)]),
            "pint_cuda_sharded_binary": (i, [i, LP, i, C.POINTER(i), C.POINTER(vp), C.POINTER(vp), C.POINTER(vp),
                                            C.POINTER(sz), C.POINTER(vp)]),
            "pint_cuda_sharded_fused_unpack_arith_repack": (i, [i, LP, i, C.POINTER(i), C.POINTER(vp), C.POINTER(vp),
                                                               C.POINTER(vp), C.POINTER(sz), C.POINTER(vp)]),
            "pint_cuda_sharded_unpack": (i, [LP, i, i, C.POINTER(i), C.POINTER(vp), C.POINTER(vp), i, C.POINTER(sz),
                                            C.POINTER(vp)]),
            "pint_cuda_sharded_repack": (i, [LP, i, C.POINTER(i), C.POINTER(vp), i, C.POINTER(vp), C.POINTER(sz),
                                            C.POINTER(vp)]),
            "pint_cuda_host_alloc": (i, [C.POINTER(vp), sz]),
            "pint_cuda_host_alloc_flags": (i, [C.POINTER(vp), sz, C.c_uint]),
            "pint_cuda_host_free": (i, [vp]),
            "pint_cuda_host_release": (i, []),
            "pint_cuda_sharded_binary_multi_host": (i, [i, C.POINTER(i), LP, i, C.POINTER(i), vp, vp, C.POINTER(vp), sz]),
            "pint_cuda_scatter_shards": (i, [vp, i, u32, sz, i, C.POINTER(i), C.POINTER(vp), C.POINTER(vp)]),
            "pint_cuda_gather_shards": (i, [vp, i, u32, sz, i, C.POINTER(i), C.POINTER(vp), C.POINTER(vp)]),
            "pint_cuda_enable_peer_access": (i, [i, C.POINTER(i)]),
            "pint_cuda_binary_host": (i, [i, LP, vp, vp, vp, sz]),
            "pint_cuda_binary_multi_host": (i, [i, C.POINTER(i), LP, vp, vp, C.POINTER(vp), sz]),
        }
        for name, (res, args) in sigs.items():
            f = getattr(L, name)
            f.restype = res
            f.argtypes = args
        L._pint_symbols = sorted(sigs)
        _lib = L
    return _lib


def error_string(rc):
    return lib().pint_cuda_error_string(rc).decode()


def check(rc, what):
    if rc != 0:
        raise PintCudaError(rc, what)
