"""ctypes binding of libsgc_b200.so (the C ABI declared in include/sgc_b200.h).

There is no fallback: if the shared library is missing or a call fails, this module raises.
Build the library with ``python -c "import __graft_entry__ as g; g.build()"`` or
``make -C sgc_b200/csrc``.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libsgc_b200.so")
# A/B builds of the same library (tools/sweep_occupancy.py): SGC_B200_LIB names another in-tree .so
if os.environ.get("SGC_B200_LIB"):
    LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), os.path.basename(os.environ["SGC_B200_LIB"]))

c_i32, c_i64, c_u32, c_sz, c_p = (ctypes.c_int32, ctypes.c_int64, ctypes.c_uint32, ctypes.c_size_t,
                                  ctypes.c_void_p)

HOP_FUSED_NORM = 1
HOP_STD_DIV = 2
HOP_STD_RCP = 4
HOP_PAD_OWNED = 8
HOP_ACCUMULATE = 16
HOP_PITCH_OWNED = 32
MAX_PEERS = 15
IPC_HANDLE_BYTES = 64


class HopArgs(ctypes.Structure):
    """struct sgc_hop_args (include/sgc_b200.h)."""
    _fields_ = [
        ("aval_d", c_p), ("dinv_d", c_p), ("dinv_rows_d", c_p), ("sval_d", c_p),
        ("x_d", c_p), ("ldx", c_i64),
        ("y_d", c_p), ("ldy", c_i64),
        ("partial_d", c_p), ("ldp", c_i64),
        ("mean_d", c_p), ("std_d", c_p),
        ("f", c_i32), ("slab_floats", c_i32), ("flags", c_u32), ("hints", c_i32),
        ("n_peers", c_i32), ("push_tile_tasks", c_i32), ("peer_y_d", c_p * 15),
        ("peer_mask_d", c_p),
    ]


# name -> (restype, argtypes); every symbol include/sgc_b200.h declares
SIGNATURES = {
    "sgc_version": (ctypes.c_char_p, []),
    "sgc_last_error": (ctypes.c_char_p, []),
    "sgc_device_check": (c_i32, [c_i32]),
    "sgc_launch_count": (c_i64, []),
    "sgc_l2_set_persisting_bytes": (c_i32, [c_i64, ctypes.POINTER(c_i64)]),
    "sgc_l2_read_probe": (c_i32, [c_p, c_i64, c_i64, c_i32, c_i32, c_i32, c_p, c_p]),
    "sgc_copy_2d": (c_i32, [c_p, c_i64, c_p, c_i64, c_i64, c_i64, c_p]),
    "sgc_host_all_equal_f64": (c_i32, [c_p, c_i64, ctypes.c_double, c_i32, c_i64]),
    "sgc_scan_workspace_bytes": (c_sz, [c_i64]),
    "sgc_scan_exclusive_i64": (c_i32, [c_p, c_i64, c_p, c_sz, c_p]),
    "sgc_csr_check_canonical": (c_i32, [c_i64, c_i64, c_p, c_p, c_p, c_p]),
    "sgc_csr_canonicalize_workspace_bytes": (c_sz, [c_i64, c_i64]),
    "sgc_csr_canonicalize": (c_i32, [c_i64, c_i64, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_sz, c_p]),
    "sgc_csr_selfloop_count": (c_i32, [c_i64, c_i64, c_p, c_p, c_p, c_p, c_p, c_sz, c_p]),
    "sgc_csr_selfloop_build": (c_i32, [c_i64, c_i64, c_p, c_p, c_p, c_p, c_p, c_p, c_p]),
    "sgc_csr_augment_count": (c_i32, [c_i64, c_i64, c_i64, c_i64, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_sz, c_p]),
    "sgc_csr_augment_build": (c_i32, [c_i64, c_i64, c_i64, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p]),
    "sgc_norm_values": (c_i32, [c_i64, c_i64, c_p, c_p, c_p, c_p, c_p, c_i32, c_p, c_p, c_p, c_p]),
    "sgc_row_normalize": (c_i32, [c_i64, c_p, c_p, c_i32, c_p, c_p, c_p, c_p]),
    "sgc_spmm_hop_sparse_x": (c_i32, [c_i64, c_p, c_p, c_p, c_p, c_p, c_p, c_i32, c_p, c_i64, c_p]),
    "sgc_gather_columns_dense": (c_i32, [c_i64, c_i64, c_p, c_p, c_p, c_p, c_p, c_i64, c_p]),
    "sgc_row_minmax": (c_i32, [c_p, c_i64, c_i64, c_i64, c_p, c_p, c_p]),
    "sgc_transpose_scale": (c_i32, [c_p, c_i64, c_i64, c_p, c_i64, c_p, c_p, c_p, c_i64, c_p]),
    "sgc_degree_inv_sqrt": (c_i32, [c_i64, c_p, c_p, c_p, c_p, c_p]),
    "sgc_norm_count": (c_i32, [c_i64, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_sz, c_p]),
    "sgc_norm_build": (c_i32, [c_i64, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p]),
    "sgc_csr_to_coo": (c_i32, [c_i64, c_p, c_p, c_i64, c_p, c_p]),
    "sgc_coo_to_csr": (c_i32, [c_i64, c_i64, c_p, c_p, c_p, c_p, c_p, c_p]),
    "sgc_col_stats_workspace_bytes": (c_sz, [c_i64, c_i64]),
    "sgc_col_stats": (c_i32, [c_p, c_i64, c_i64, c_i64, c_p, c_p, c_p, c_p, c_p, c_sz, c_p]),
    "sgc_col_stats_static": (c_i32, [c_p, c_i64, c_i64, c_i64, c_p, c_p, c_p, c_p, c_p]),
    "sgc_standardize_inplace": (c_i32, [c_p, c_i64, c_i64, c_i64, c_p, c_p, c_p]),
    "sgc_spmm_plan_workspace_bytes": (c_sz, [c_i64, c_i64, c_i32]),
    "sgc_spmm_plan_create": (c_i32, [c_i64, c_i64, c_i64, c_p, c_p, c_i32, c_p, c_sz,
                                     ctypes.POINTER(c_p), c_p]),
    "sgc_spmm_plan_destroy": (None, [c_p]),
    "sgc_spmm_plan_partial_rows": (c_i64, [c_p]),
    "sgc_spmm_hot_workspace_bytes": (c_sz, [c_i64]),
    "sgc_spmm_plan_tag_hot": (c_i32, [c_p, c_i64, c_p, c_p, c_sz, ctypes.POINTER(c_i64), c_p]),
    "sgc_spmm_hop": (c_i32, [c_p, ctypes.POINTER(HopArgs), c_p]),
    "sgc_precompute": (c_i32, [c_p, ctypes.POINTER(HopArgs), c_i32, c_p, c_p, c_i64, c_p]),
    "sgc_exp_spmm_hop_bulk_gather": (c_i32, [c_p, ctypes.POINTER(HopArgs), c_i32, c_i32, c_p]),
    "sgc_exp_spmm_hop_gather4": (c_i32, [c_p, ctypes.POINTER(HopArgs), c_i32, c_i32, c_p]),
    "sgc_sort_workspace_bytes": (c_sz, [c_i64]),
    "sgc_sort_by_key_i32": (c_i32, [c_i64, c_i32, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_sz, c_p]),
    "sgc_csr_expand_rows": (c_i32, [c_i64, c_p, c_i32, c_p, c_p]),
    "sgc_coo_reduce_workspace_bytes": (c_sz, [c_i64]),
    "sgc_coo_reduce_to_csr": (c_i32, [c_i64, c_i64, c_p, c_p, c_p, c_i32, c_i32, c_p, c_p, c_p, c_p, c_sz, c_p]),
    "sgc_subgraph_colmap": (c_i32, [c_i64, c_i64, c_p, c_p, c_p]),
    "sgc_subgraph_count": (c_i32, [c_i64, c_p, c_p, c_p, c_p, c_p, c_p, c_sz, c_p]),
    "sgc_subgraph_fill": (c_i32, [c_i64, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p, c_p]),
    "sgc_symm_alloc": (c_i32, [c_sz, ctypes.POINTER(c_p)]),
    "sgc_symm_free": (c_i32, [c_p]),
    "sgc_ipc_export": (c_i32, [c_p, c_p]),
    "sgc_ipc_open": (c_i32, [c_p, ctypes.POINTER(c_p)]),
    "sgc_ipc_close": (c_i32, [c_p]),
    "sgc_push_bytes_bulk": (c_i32, [c_p, c_i64, ctypes.POINTER(c_p), c_i32, c_i32, c_i32, c_p]),
    "sgc_push_rows": (c_i32, [c_p, c_i64, c_i64, c_i32, ctypes.POINTER(c_p), c_i32, c_i64, c_p]),
}

_lib = None


class SgcError(RuntimeError):
    pass


def lib():
    """The loaded library; raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise SgcError(
                "libsgc_b200.so is not built (%s missing); run `make -C sgc_b200/csrc` -- there is "
                "no CPU or PyTorch fallback for the propagation path" % LIB_PATH)
        handle = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)   # AttributeError if the symbol is not exported
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def check(status, what=""):
    if status != 0:
        msg = lib().sgc_last_error()
        raise SgcError("%s failed with status %d: %s" % (what or "libsgc_b200 call", status,
                                                         msg.decode() if msg else "?"))


def call(name, *args):
    """Invoke an int32-status entry point and raise on failure."""
    check(getattr(lib(), name)(*args), name)


def ptr(t):
    """Device pointer of a torch tensor (None -> NULL)."""
    return None if t is None else t.data_ptr()


def launch_count():
    return int(lib().sgc_launch_count())
