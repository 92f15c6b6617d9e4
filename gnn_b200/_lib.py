"""ctypes binding of libgnn_b200.so (the C ABI declared in include/gnnb.h).

The library is built in-tree by ``gnn_b200/csrc/Makefile`` (``__graft_entry__.build()``) into
``gnn_b200/lib/libgnn_b200.so``.  There is no CPU fallback anywhere in this package: if the library is
missing, or the current device is not an sm_100 GPU, every operator raises.
"""
import ctypes
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("GNNB_LIB") or os.path.join(_HERE, "lib", "libgnn_b200.so")   # GNNB_LIB: a differently tuned build
CSRC_DIR = os.path.join(_HERE, "csrc")

_c = ctypes
_i64, _i32, _sz, _vp, _u64 = _c.c_int64, _c.c_int, _c.c_size_t, _c.c_void_p, _c.c_uint64
_psz = _c.POINTER(_c.c_size_t)

# name -> (restype, argtypes); mirrors include/gnnb.h one to one (tests check the two against each other)
PROTOTYPES = {
    "gnnb_version": (_i32, []),
    "gnnb_last_error": (_c.c_char_p, []),
    "gnnb_check_device": (_i32, []),
    "gnnb_synth_genotypes": (_i32, [_vp, _i64, _i64, _i64, _i64, _i64, _u64, _u64, _vp]),
    "gnnb_gather_rows_i8": (_i32, [_vp, _i64, _vp, _i64, _i64, _vp, _i64, _vp]),
    "gnnb_i8_to_f32": (_i32, [_vp, _i64, _i64, _i64, _vp, _i64, _vp]),
    "gnnb_row_sqnorm_i8": (_i32, [_vp, _i64, _i64, _i64, _vp, _vp]),
    "gnnb_knn_workspace_bytes": (_i32, [_i64, _i64, _i64, _i32, _i32, _psz]),
    "gnnb_knn_i8": (_i32, [_vp, _i64, _vp, _i64, _i64, _vp, _i64, _vp, _i64, _i64, _i64, _vp, _vp, _i32,
                           _i32, _vp, _vp, _i32, _i32, _vp, _sz, _vp]),
    "gnnb_pack_fp4": (_i32, [_vp, _i64, _i64, _i64, _vp, _i64, _vp, _vp]),
    "gnnb_knn_fp4": (_i32, [_vp, _i64, _vp, _i64, _i64, _vp, _i64, _vp, _i64, _i64, _i64, _vp, _vp, _i32,
                            _i32, _vp, _vp, _i32, _vp, _sz, _vp]),
    "gnnb_knn_f64": (_i32, [_vp, _i64, _i64, _i64, _vp, _i64, _i64, _i64, _i64, _i32, _vp, _vp, _i32, _vp, _sz, _vp]),
    "gnnb_knn_insert_self": (_i32, [_vp, _vp, _i64, _i32, _i64, _vp]),
    "gnnb_unpack_2bit": (_i32, [_vp, _i64, _i64, _i64, _vp, _i64, _vp]),
    "gnnb_base3_to_2bit": (_i32, [_vp, _i64, _i64, _i64, _vp, _i64, _vp, _vp]),
    "gnnb_pack_2bit": (_i32, [_vp, _i64, _i64, _i64, _vp, _i64, _vp, _vp]),
    "gnnb_scale_columns_i8": (_i32, [_vp, _i64, _i64, _i64, _vp, _i64, _vp, _i64, _vp]),
    "gnnb_graph_workspace_bytes": (_i32, [_i64, _i32, _psz]),
    "gnnb_graph_count": (_i32, [_vp, _vp, _i64, _i32, _i32, _vp, _vp, _sz, _vp]),
    "gnnb_graph_fill": (_i32, [_i64, _i32, _i32, _i64, _vp, _vp, _vp, _sz, _vp]),
    "gnnb_graph_knn_edges": (_i32, [_vp, _vp, _i64, _i32, _i32, _i32, _vp, _vp, _vp, _sz, _vp]),
    "gnnb_csr_workspace_bytes": (_i32, [_i64, _i64, _psz]),
    "gnnb_csr_count": (_i32, [_vp, _i64, _i64, _i32, _vp, _vp, _sz, _vp]),
    "gnnb_csr_build": (_i32, [_vp, _vp, _i64, _i64, _i32, _i64, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "gnnb_spmm": (_i32, [_vp, _vp, _vp, _i64, _vp, _i64, _i64, _vp, _vp, _i64, _vp, _vp, _vp, _vp, _i64, _vp, _sz, _vp]),
    "gnnb_spmm_long_row": (_i32, []),
    "gnnb_spmm_workspace_bytes": (_i32, [_i64, _i64, _psz]),
    "gnnb_linear_workspace_bytes": (_i32, [_i64, _i64, _i64, _i32, _psz]),
    "gnnb_linear_fwd": (_i32, [_vp, _i32, _i64, _i64, _i64, _vp, _i64, _i64, _vp, _i64, _i32, _vp, _sz, _vp]),
    "gnnb_linear_bwd_w": (_i32, [_vp, _i32, _i64, _i64, _i64, _vp, _i64, _i64, _vp, _i64, _i32, _vp, _sz, _vp]),
    "gnnb_linear_bwd_x": (_i32, [_vp, _i64, _i64, _i64, _vp, _i64, _i64, _vp, _i64, _vp]),
    "gnnb_linear_supported": (_i32, [_vp, _i32, _i64, _i64, _i64, _i64, _i32]),
    "gnnb_linear_bwd_w_wave_markers": (_i64, [_i32]),
    "gnnb_linear_prepare_dh": (_i32, [_i32, _i64, _vp, _i64, _i64, _vp, _sz, _vp]),
    "gnnb_linear_bwd_w_prepared": (_i32, [_vp, _i32, _i64, _i64, _i64, _i64, _vp, _i64, _vp, _sz, _vp]),
    "gnnb_colsum_workspace_bytes": (_i32, [_i64, _psz]),
    "gnnb_colsum": (_i32, [_vp, _i64, _i64, _i64, _vp, _vp, _sz, _vp]),
    "gnnb_colsum_static": (_i32, [_vp, _i64, _i64, _i64, _vp, _vp]),
    "gnnb_head_supported": (_i32, [_i64, _i64]),
    "gnnb_head_workspace_bytes": (_i32, [_i64, _i64, _i64, _psz]),
    "gnnb_head_fwd": (_i32, [_vp, _i64, _i64, _i64, _vp, _i64, _i64, _vp, _vp, _i64, _vp]),
    "gnnb_head_bwd": (_i32, [_vp, _i64, _vp, _i64, _i64, _i64, _vp, _i64, _i64, _vp, _i64, _vp, _i64, _vp, _vp, _sz, _vp]),
    "gnnb_head_dropout_fwd": (_i32, [_vp, _i64, _i64, _i64, _vp, _i64, _i64, _vp, _vp, _i64, _c.c_double, _vp, _vp, _vp]),
    "gnnb_head_dropout_bwd": (_i32, [_vp, _i64, _vp, _i64, _i64, _i64, _vp, _i64, _i64, _vp, _i64, _vp, _i64, _vp, _c.c_double, _vp,
                                     _vp, _sz, _vp]),
    "gnnb_adam_step": (_i32, [_i32, _vp, _vp, _vp, _vp, _vp, _vp, _c.c_float, _c.c_float, _c.c_float, _c.c_float, _vp]),
    "gnnb_allreduce_mean_multimem": (_i32, [_vp, _i64, _i32, _i32, _i32, _vp]),
    "gnnb_allreduce_mean_p2p": (_i32, [_vp, _i64, _i32, _i32, _i32, _vp]),
    "gnnb_first_occurrence": (_i32, [_vp, _i64, _i64, _vp, _vp]),
    "gnnb_partition_walk_host": (_i64, [_vp, _i64, _i64, _vp, _i64, _vp, _vp, _vp]),
}

_lib = None


class GnnbError(RuntimeError):
    pass


def build(verbose=False):
    """Compile libgnn_b200.so for sm_100a with nvcc (cross-compiles without a GPU)."""
    res = subprocess.run(["make", "-C", CSRC_DIR, "-j8"], capture_output=True, text=True)
    if verbose or res.returncode != 0:
        print(res.stdout[-4000:])
        print(res.stderr[-4000:])
    if res.returncode != 0:
        raise GnnbError("building libgnn_b200.so failed (see output above)")
    return LIB_PATH


def lib():
    """The loaded library with prototypes set.  Raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise GnnbError(
                f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                "(gnn_b200 has no CPU fallback)")
        handle = ctypes.CDLL(LIB_PATH)
        for name, (restype, argtypes) in PROTOTYPES.items():
            fn = getattr(handle, name)
            fn.restype = restype
            fn.argtypes = argtypes
        _lib = handle
    return _lib


def check(rc, what=""):
    if rc < 0:
        msg = lib().gnnb_last_error().decode("utf-8", "replace")
        raise GnnbError(f"{what}: {msg} (code {rc})" if what else f"{msg} (code {rc})")
    return rc


def call(name, *args):
    """Invoke an int-returning entry point; raise GnnbError with the library's message on failure."""
    return check(getattr(lib(), name)(*args), name)


def workspace_bytes(name, *args):
    out = ctypes.c_size_t(0)
    call(name, *args, ctypes.byref(out))
    return int(out.value)
