"""ctypes loader for libb200spmv.so (the C ABI of include/b200spmv.h)."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B200SPMV_LIB") or os.path.join(_HERE, "lib", "libb200spmv.so")


class B200Error(RuntimeError):
    pass


_lib = None

_p = C.c_void_p
_i32, _i64, _f64, _f32, _int, _sz = C.c_int32, C.c_int64, C.c_double, C.c_float, C.c_int, C.c_size_t

# name -> argtypes; every function returns int unless noted
SIGNATURES = {
    "b200_abi_version": [],
    "b200_device_count": [C.POINTER(_int)],
    "b200_set_device": [_int],
    "b200_get_num_sms": [C.POINTER(_int)],
    "b200_malloc": [C.POINTER(_p), _sz],
    "b200_free": [_p],
    "b200_memset_async": [_p, _int, _sz, _p],
    "b200_memcpy_d2d_async": [_p, _p, _sz, _p],
    "b200_memcpy_h2d_async": [_p, _p, _sz, _p],
    "b200_memcpy_d2h_async": [_p, _p, _sz, _p],
    "b200_host_alloc_pinned": [C.POINTER(_p), _sz],
    "b200_host_free_pinned": [_p],
    "b200_stream_create": [C.POINTER(_p)],
    "b200_stream_create_priority": [C.POINTER(_p), _int],
    "b200_ipc_alloc": [C.POINTER(_p), _sz, C.c_char_p],
    "b200_ipc_open": [C.POINTER(_p), C.c_char_p],
    "b200_ipc_close": [_p],
    "b200_stream_write_u32": [_p, _p, C.c_uint32],
    "b200_stream_wait_u32_geq": [_p, _p, C.c_uint32],
    "b200_stream_destroy": [_p],
    "b200_stream_synchronize": [_p],
    "b200_event_create": [C.POINTER(_p), _int],
    "b200_event_destroy": [_p],
    "b200_event_record": [_p, _p],
    "b200_event_synchronize": [_p],
    "b200_stream_wait_event": [_p, _p],
    "b200_event_elapsed_ms": [_p, _p, C.POINTER(_f32)],
    "b200_csr_plan_create": [C.POINTER(_p), _i32, _i32, _i64, _p, _int, _p],
    "b200_csr_plan_destroy": [_p],
    "b200_csr_plan_info": [_p, C.POINTER(_i32), C.POINTER(_i32), C.POINTER(_i64)],
    "b200_csr_plan_tune": [_p, _p, _int, _int, _int, _int, _int, _p],
    "b200_csr_plan_chunking": [_p, _p, _i32, _p, _p, _p, _p, _p],
    "b200_csr_spmv_tiles_f64": [_p, _p, _p, _p, _f64, _p, _f64, _p, _i32, _i32, _p],
    "b200_csr_symv_tiles_f64": [_p, _p, _p, _p, _p, _f64, _p, _p, _i32, _i32, _p],
    "b200_csr_spmv_f64": [_p, _p, _p, _p, _f64, _p, _f64, _p, _p],
    "b200_csr_spmv_f32": [_p, _p, _p, _p, _f32, _p, _f32, _p, _p],
    "b200_csr_spmv_dot_f64": [_p, _p, _p, _p, _p, _p, _p, _p],
    "b200_csr_symv_dot_f64": [_p, _p, _p, _p, _p, _p, _p, _p, _p],
    "b200_csr_plan_rowlist_dot_ok": [_p, _p],
    "b200_csr_spmv_rowlist_dot_f64": [_p, _p, _p, _p, _p, _p, _p, _p, _p],
    "b200_csr_spmv_rowlist_dot_halo_f64": [_p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _i32, _p],
    "b200_csr_symv_f64": [_p, _p, _p, _p, _p, _f64, _p, _f64, _p, _p],
    "b200_csr_symv_f32": [_p, _p, _p, _p, _p, _f32, _p, _f32, _p, _p],
    "b200_gather_f64": [_i32, _p, _p, _p, _p],
    "b200_gather_f32": [_i32, _p, _p, _p, _p],
    "b200_scatter_add_f64": [_i32, _p, _p, _p, _p],
    "b200_scatter_add_f32": [_i32, _p, _p, _p, _p],
    "b200_nccl_unique_id": [C.c_char_p],
    "b200_nccl_comm_create": [C.POINTER(_p), _int, _int, C.c_char_p],
    "b200_nccl_comm_destroy": [_p],
    "b200_halo_exchange": [_p, _int, _p, _p, _p, _p, _p, _p, _p, _int, _p],
    "b200_allreduce_sum_f64": [_p, _p, _i32, _p],
    "b200_cg_dot_f64": [_i32, _p, _p, _p, _p, _p, _p],
    "b200_cg_dot_peer_f64": [_i32, _p, _p, _p, _p, _p, _p, _p],
    "b200_cg_dot_static_f64": [_i32, _p, _p, _p, _p],
    "b200_peer_reduce_mailbox_bytes": [_i32, _p],
    "b200_peer_reduce_create": [_p, _i32, _i32, _p, _p],
    "b200_peer_reduce_destroy": [_p],
    "b200_peer_allreduce_sum_f64": [_p, _p, _i32, _p],
    "b200_cg_init_state": [_p, _p, _f64, _i32, _p],
    "b200_cg_fused_step_f64": [_i32, _p, _p, _p, _p, _p, _p, _i32, _p, _int, _p],
    "b200_cg_fused_step_halo_f64": [_i32, _p, _p, _p, _p, _p, _p, _i32, _p, _int, _p, _i32, _p, _p],
    "b200_graph_capture_begin": [_p],
    "b200_graph_capture_end": [_p, C.POINTER(_p)],
    "b200_graph_launch": [_p, _p],
    "b200_graph_destroy": [_p],
    "b200_halo_desc_create": [C.POINTER(_p), _i32, _p, _p, _p, _p, _p, _p, _p, _p, _i64, _p],
    "b200_halo_desc_destroy": [_p],
    "b200_halo_push_f64": [_p, _i32, _p, _p, _p],
    "b200_halo_push_f32": [_p, _i32, _p, _p, _p],
    "b200_halo_wait_unpack_f64": [_p, _i64, _p, _p],
    "b200_halo_wait_unpack_f32": [_p, _i64, _p, _p],
    "b200_cg_update_xr_f64": [_i32, _p, _p, _p, _p, _p, _p],
    "b200_cg_update_xr_peer_f64": [_i32, _p, _p, _p, _p, _p, _p, _p],
    "b200_cg_update_p_f64": [_i32, _p, _p, _p, _p, _i32, _p],
    "b200_gen_laplace3d_rowptr": [_int, _int, _i64, _i64, _i64, _int, _int, _p, _p,
                                  C.POINTER(_i64), C.POINTER(_i64), _p],
    "b200_gen_laplace3d_fill": [_int, _int, _i64, _i64, _i64, _int, _int, _p, _p, _p, _p, _p, _p, _p, _p],
}

CG_STATE_DOUBLES = 4096
DOT_SCRATCH_DOUBLES = 4096
NCCL_UNIQUE_ID_BYTES = 128


def lib():
    """Load the CUDA library; raise (never fall back) if it is not there."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise B200Error(
                f"{LIB_PATH} not found: the CUDA extension is not built "
                "(run `python -c 'import __graft_entry__ as g; g.build()'`). "
                "There is no CPU fallback in this package."
            )
        L = C.CDLL(LIB_PATH)
        for name, args in SIGNATURES.items():
            f = getattr(L, name)
            f.argtypes = args
            f.restype = C.c_int
        L.b200_last_error.argtypes = []
        L.b200_last_error.restype = C.c_char_p
        _lib = L
    return _lib


def check(rc):
    if rc != 0:
        raise B200Error(f"b200spmv error {rc}: {lib().b200_last_error().decode(errors='replace')}")
