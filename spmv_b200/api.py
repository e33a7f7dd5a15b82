"""Python mirror of the reference's API for the hot path, over libspmv_b200.so.

Same names and argument meaning as the reference's C++ classes (spmv::L2GMap,
spmv::Matrix<double>, spmv::cg, read_petsc_binary_matrix, create_A, the executors), so
the parity tests read like the reference's own tests/test_spmv.cpp.  All work happens
in the C++ host layer / CUDA kernels; this file only marshals arguments (ctypes).
PyTorch appears only as plumbing: device tensors carry executor memory, and
torch.distributed carries the SETUP collectives of a multi-process run.
"""
import ctypes as C
import os
import threading

import numpy as np

from ._lib import B200Error, LIB_PATH, lib as _kernel_lib

_HERE = os.path.dirname(os.path.abspath(__file__))
HOST_LIB_PATH = os.path.join(_HERE, "lib", "libspmv_b200.so")

# CommunicationModel (spmv/mpi_utils.h:43-52)
P2P_BLOCKING, P2P_NONBLOCKING, COLLECTIVE_BLOCKING, COLLECTIVE_NONBLOCKING = 0, 1, 2, 3
ONESIDED_PUT_ACTIVE, ONESIDED_PUT_PASSIVE, SHMEM, SHMEM_NODUP = 4, 5, 6, 7

_p, _i32, _i64, _f64, _int, _sz = C.c_void_p, C.c_int32, C.c_int64, C.c_double, C.c_int, C.c_size_t
_pp = C.POINTER(C.c_void_p)
BARRIER_FN = C.CFUNCTYPE(C.c_int, C.c_void_p)
ALLTOALLV_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p, C.POINTER(_i64), C.POINTER(_i64), C.c_void_p,
                           C.POINTER(_i64), C.POINTER(_i64))
RANK_FN = C.CFUNCTYPE(None, C.c_void_p)

_SIG = {
    "spmvh_exec_create_host": [_pp],
    "spmvh_exec_create_b200": [_pp, _int],
    "spmvh_exec_destroy": [_p],
    "spmvh_exec_set_stream": [_p, _p],
    "spmvh_exec_synchronize": [_p],
    "spmvh_exec_alloc": [_p, _sz, _pp],
    "spmvh_exec_free": [_p, _p],
    "spmvh_exec_copy_from_host": [_p, _p, _p, _sz],
    "spmvh_exec_copy_to_host": [_p, _p, _p, _sz],
    "spmvh_l2g_create": [_pp, _p, _i64, _p, _i64, _int],
    "spmvh_l2g_destroy": [_p],
    "spmvh_l2g_sizes": [_p, C.POINTER(_i64)],
    "spmvh_l2g_get": [_p, _int, _p, _i64, C.POINTER(_i64)],
    "spmvh_l2g_global_to_local": [_p, _i64, C.POINTER(_i32)],
    "spmvh_l2g_update_f64": [_p, _p],
    "spmvh_l2g_update_finalise_f64": [_p, _p],
    "spmvh_l2g_reverse_update_f64": [_p, _p],
    "spmvh_matrix_create": [_pp, _p, _p, _p, _p, _i64, _i64, _p, _i64, _p, _i64, _int, _int],
    "spmvh_matrix_read_petsc": [_pp, _p, C.c_char_p, _int, _int],
    "spmvh_matrix_create_A": [_pp, _p, _int],
    "spmvh_matrix_create_laplace3d": [_pp, _p, _int, _int, _i64, _int, _int, _int],
    "spmvh_matrix_destroy": [_p],
    "spmvh_matrix_info": [_p, C.POINTER(_i64)],
    "spmvh_matrix_col_map": [_p, _pp],
    "spmvh_matrix_row_map": [_p, _pp],
    "spmvh_matrix_block_dims": [_p, _int, C.POINTER(_i64)],
    "spmvh_matrix_block_copy": [_p, _int, _p, _p, _p, _p],
    "spmvh_matrix_mult_f64": [_p, _p, _p],
    "spmvh_matrix_mult_host_f64": [_p, _p, _p],
    "spmvh_l2g_allreduce_f64": [_p, _p, _i32, C.c_int],
    "spmvh_matrix_tune": [_p, _int, _int, _int, _int, _int, _int],
    "spmvh_cg": [_p, _p, _p, _p, _int, _f64, _p, _int, C.POINTER(_int)],
    "spmvh_cg_phases": [_p, _p, _p, _p, _int, _f64, C.POINTER(_f64), C.POINTER(_int)],
    "spmvh_l2g_halo_mode": [_p, C.c_char_p, _int],
    "spmvh_l2g_update_f32": [_p, _p],
    "spmvh_l2g_update_finalise_f32": [_p, _p],
    "spmvh_l2g_reverse_update_f32": [_p, _p],
    "spmvh_matrix_create_f32": [_pp, _p, _p, _p, _p, _i64, _i64, _p, _i64, _p, _i64, _int, _int],
    "spmvh_matrix_destroy_f32": [_p],
    "spmvh_matrix_info_f32": [_p, C.POINTER(_i64)],
    "spmvh_matrix_col_map_f32": [_p, _pp],
    "spmvh_matrix_mult_f32": [_p, _p, _p],
    "spmvh_matrix_mult_host_f32": [_p, _p, _p],
    "spmvh_write_petsc_matrix": [C.c_char_p, _i32, _i32, _p, _p, _p],
    "spmvh_write_petsc_vector": [C.c_char_p, _i32, _p],
    "spmvh_read_petsc_vector": [_p, C.c_char_p, C.POINTER(C.c_void_p), C.POINTER(_i64)],
    "spmvh_powerlaw_counts": [_i64, C.c_uint64, _f64, _f64, _i64, _i64, _f64, _i64, _i64, _p],
    "spmvh_powerlaw_fill": [_i64, C.c_uint64, _f64, _f64, _i64, _i64, _f64, _i64, _i64, _p, _p, _p],
    "spmvh_write_powerlaw_petsc": [C.c_char_p, _i64, C.c_uint64, _f64, _f64, _i64, _i64, _f64, C.POINTER(_i64)],
    "spmv_shim_run_ranks": [_int, RANK_FN, _p],
    "spmv_shim_set_world_callbacks": [_int, _int, BARRIER_FN, ALLTOALLV_FN, _p],
}

_host = None


def host_lib():
    """Load the C++ host layer (which links the CUDA kernel library); no fallback."""
    global _host
    if _host is None:
        _kernel_lib()  # loads libb200spmv.so first so the dependency resolves in-tree
        if not os.path.exists(HOST_LIB_PATH):
            raise B200Error(f"{HOST_LIB_PATH} not found: run __graft_entry__.build()")
        H = C.CDLL(HOST_LIB_PATH, mode=C.RTLD_GLOBAL)
        for name, args in _SIG.items():
            f = getattr(H, name)
            f.argtypes = args
            f.restype = C.c_int
        H.spmvh_last_error.argtypes = []
        H.spmvh_last_error.restype = C.c_char_p
        _host = H
    return _host


def _ck(rc):
    if rc != 0:
        raise B200Error(host_lib().spmvh_last_error().decode(errors="replace"))


def _np_ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _dev_ptr(x):
    """Device pointer of a torch tensor, or an int address."""
    if x is None:
        return None
    if hasattr(x, "data_ptr"):
        return C.c_void_p(x.data_ptr())
    return C.c_void_p(int(x))


# ----------------------------------------------------------------------------
# worlds: thread-ranks in one process, or one process per GPU over torch.distributed
# ----------------------------------------------------------------------------
def run_ranks(n, fn):
    """Run fn(rank) on n thread-ranks (spmv_shim_run_ranks); MPI_COMM_WORLD inside is the
    n-rank world.  Exceptions raised by fn are re-raised here."""
    errors, counter, lock = [], [0], threading.Lock()

    def body(_arg):
        try:
            fn(world_rank())
        except BaseException as e:  # noqa: BLE001
            with lock:
                errors.append(e)

    del counter
    cb = RANK_FN(body)
    rc = host_lib().spmv_shim_run_ranks(n, cb, None)
    if errors:
        raise errors[0]
    if rc != 0:
        raise B200Error("a thread-rank failed")


_world_keepalive = []


def init_world_from_torch(group=None):
    """Install torch.distributed (any backend with all_gather; gloo for host tensors) as
    the transport of MPI_COMM_WORLD for this process: only setup traffic goes through it."""
    import torch
    import torch.distributed as dist

    rank, size = dist.get_rank(group), dist.get_world_size(group)
    backend = dist.get_backend(group)
    dev = torch.device("cuda", torch.cuda.current_device()) if backend == "nccl" else torch.device("cpu")

    def barrier(_ctx):
        try:
            dist.barrier(group)
            return 0
        except Exception:  # noqa: BLE001
            return 1

    def alltoallv(_ctx, sbuf, scnt, sdsp, rbuf, rcnt, rdsp):
        try:
            sc = [int(scnt[i]) for i in range(size)]
            sd = [int(sdsp[i]) for i in range(size)]
            rc_ = [int(rcnt[i]) for i in range(size)]
            rd = [int(rdsp[i]) for i in range(size)]
            # pack my outgoing bytes rank by rank, exchange everything with all_gather
            # (setup path: simple and backend-independent), then pick my pieces
            out = bytearray()
            for i in range(size):
                if sc[i]:
                    out += C.string_at(sbuf + sd[i], sc[i])
            lens = torch.tensor(sc, dtype=torch.int64, device=dev)
            all_lens = [torch.zeros(size, dtype=torch.int64, device=dev) for _ in range(size)]
            dist.all_gather(all_lens, lens, group=group)
            all_lens = [t.cpu().tolist() for t in all_lens]
            totals = [sum(l) for l in all_lens]
            mx = max(max(totals), 1)
            payload = torch.zeros(mx, dtype=torch.uint8)
            if len(out):
                payload[: len(out)] = torch.frombuffer(out, dtype=torch.uint8)
            payload = payload.to(dev)
            gathered = [torch.zeros(mx, dtype=torch.uint8, device=dev) for _ in range(size)]
            dist.all_gather(gathered, payload, group=group)
            for src in range(size):
                n = all_lens[src][rank]
                if n != rc_[src]:
                    return 2
                if n:
                    off = sum(all_lens[src][:rank])
                    piece = gathered[src][off:off + n].cpu().numpy().tobytes()
                    C.memmove(rbuf + rd[src], piece, n)
            return 0
        except Exception:  # noqa: BLE001
            import traceback
            traceback.print_exc()
            return 1

    b, a = BARRIER_FN(barrier), ALLTOALLV_FN(alltoallv)
    _world_keepalive.append((b, a))
    rc = host_lib().spmv_shim_set_world_callbacks(rank, size, b, a, None)
    if rc != 0:
        raise B200Error("could not install the torch.distributed world")
    return rank, size


def _world():
    H = host_lib()
    H.spmv_shim_comm_world.restype = C.c_void_p
    H.spmv_shim_comm_world.argtypes = []
    H.MPI_Comm_rank.argtypes = [C.c_void_p, C.POINTER(C.c_int)]
    H.MPI_Comm_size.argtypes = [C.c_void_p, C.POINTER(C.c_int)]
    return H, H.spmv_shim_comm_world()


def world_rank():
    H, w = _world()
    r = C.c_int()
    H.MPI_Comm_rank(w, C.byref(r))
    return r.value


def world_size():
    H, w = _world()
    s = C.c_int()
    H.MPI_Comm_size(w, C.byref(s))
    return s.value


# ----------------------------------------------------------------------------
# executors
# ----------------------------------------------------------------------------
class Executor:
    """spmv::DeviceExecutor handle.  `ReferenceExecutor.create()` is host memory only
    (no compute); `B200Executor.create(device_id)` is the compute backend."""

    def __init__(self, handle, kind):
        self.h, self.kind = handle, kind

    def set_cuda_stream(self, stream):
        _ck(host_lib().spmvh_exec_set_stream(self.h, C.c_void_p(int(stream))))

    def use_torch_stream(self):
        import torch
        self.set_cuda_stream(torch.cuda.current_stream().cuda_stream)

    def synchronize(self):
        _ck(host_lib().spmvh_exec_synchronize(self.h))

    def alloc(self, num_bytes):
        p = C.c_void_p()
        _ck(host_lib().spmvh_exec_alloc(self.h, num_bytes, C.byref(p)))
        return p.value

    def free(self, ptr):
        _ck(host_lib().spmvh_exec_free(self.h, C.c_void_p(ptr)))

    def copy_from_host(self, d_ptr, array):
        a = np.ascontiguousarray(array)
        _ck(host_lib().spmvh_exec_copy_from_host(self.h, _dev_ptr(d_ptr), _np_ptr(a), a.nbytes))

    def copy_to_host(self, array, d_ptr):
        _ck(host_lib().spmvh_exec_copy_to_host(self.h, _np_ptr(array), _dev_ptr(d_ptr), array.nbytes))

    def close(self):
        if self.h:
            host_lib().spmvh_exec_destroy(self.h)
            self.h = None


class ReferenceExecutor(Executor):
    @staticmethod
    def create():
        h = C.c_void_p()
        _ck(host_lib().spmvh_exec_create_host(C.byref(h)))
        return ReferenceExecutor(h, "cpu")


class B200Executor(Executor):
    @staticmethod
    def create(device_id=0):
        h = C.c_void_p()
        _ck(host_lib().spmvh_exec_create_b200(C.byref(h), int(device_id)))
        return B200Executor(h, "gpu")


CudaExecutor = B200Executor


# ----------------------------------------------------------------------------
# L2GMap
# ----------------------------------------------------------------------------
class L2GMap:
    """spmv::L2GMap(comm = MPI_COMM_WORLD, local_size, ghosts, exec, cm)."""

    _GET = {"ranges": (0, np.int64), "ghosts": (1, np.int64), "neighbours": (2, np.int32),
            "send_count": (3, np.int32), "recv_count": (4, np.int32), "send_offset": (5, np.int32),
            "recv_offset": (6, np.int32), "indexbuf": (7, np.int32)}

    def __init__(self, local_size=None, ghosts=None, exec=None, cm=COLLECTIVE_BLOCKING, _handle=None):
        if _handle is not None:
            self.h = _handle
            return
        g = np.ascontiguousarray(ghosts if ghosts is not None else [], dtype=np.int64)
        h = C.c_void_p()
        _ck(host_lib().spmvh_l2g_create(C.byref(h), exec.h, int(local_size), _np_ptr(g), len(g), int(cm)))
        self.h = h

    def _sizes(self):
        out = (_i64 * 8)()
        _ck(host_lib().spmvh_l2g_sizes(self.h, out))
        return list(out)

    def local_size(self):
        return self._sizes()[0]

    def num_ghosts(self):
        return self._sizes()[1]

    def global_size(self):
        return self._sizes()[2]

    def global_offset(self):
        return self._sizes()[3]

    def rank(self):
        return self._sizes()[6]

    def overlapping(self):
        return bool(self._sizes()[7])

    def global_to_local(self, g):
        out = _i32()
        _ck(host_lib().spmvh_l2g_global_to_local(self.h, int(g), C.byref(out)))
        return out.value

    def get(self, name):
        what, dt = self._GET[name]
        n = _i64()
        _ck(host_lib().spmvh_l2g_get(self.h, what, None, 0, C.byref(n)))
        out = np.zeros(n.value, dtype=dt)
        _ck(host_lib().spmvh_l2g_get(self.h, what, _np_ptr(out), n.value, C.byref(n)))
        return out

    def plan(self):
        """Every integer artefact of SURVEY.md Appendix A, as numpy arrays."""
        d = {k: self.get(k) for k in self._GET}
        s = self._sizes()
        d.update(local_size=s[0], num_ghosts=s[1], global_size=s[2], global_offset=s[3])
        return d

    @staticmethod
    def _sfx(vec):
        """L2GMap::update<T> is a template: the tensor's dtype picks the instantiation."""
        return "f32" if str(getattr(vec, "dtype", "")).endswith("float32") else "f64"

    def update(self, vec):
        _ck(getattr(host_lib(), "spmvh_l2g_update_" + self._sfx(vec))(self.h, _dev_ptr(vec)))

    def update_finalise(self, vec):
        _ck(getattr(host_lib(), "spmvh_l2g_update_finalise_" + self._sfx(vec))(self.h, _dev_ptr(vec)))

    def reverse_update(self, vec):
        _ck(getattr(host_lib(), "spmvh_l2g_reverse_update_" + self._sfx(vec))(self.h, _dev_ptr(vec)))

    def halo_mode(self):
        """'<exchange update() uses>/<exchange spmv::cg switches to>', e.g. 'copy/kernel', 'none/none'."""
        buf = C.create_string_buffer(32)
        _ck(host_lib().spmvh_l2g_halo_mode(self.h, buf, 32))
        return buf.value.decode()

    def allreduce_sum(self, vec, nccl=False):
        """Sum the (<= 4) doubles of `vec` over the ranks, in place, on the executor's stream."""
        _ck(host_lib().spmvh_l2g_allreduce_f64(self.h, _dev_ptr(vec), int(vec.numel()), int(bool(nccl))))

    def close(self):
        if self.h:
            host_lib().spmvh_l2g_destroy(self.h)
            self.h = None


# ----------------------------------------------------------------------------
# Matrix<double>
# ----------------------------------------------------------------------------
class Matrix:
    def __init__(self, handle, exec):
        self.h, self.exec = handle, exec
        self._col = self._row = None

    @staticmethod
    def create_matrix(exec, rowptr, colind, values, nrows_local, ncols_local, row_ghosts, col_ghosts,
                      symmetric=False, cm=COLLECTIVE_BLOCKING):
        """Matrix<double>::create_matrix (raw-pointer overload, Matrix.h:106-111)."""
        rp = np.ascontiguousarray(rowptr, dtype=np.int32)
        ci = np.ascontiguousarray(colind, dtype=np.int32)
        v = np.ascontiguousarray(values, dtype=np.float64)
        rg = np.ascontiguousarray(row_ghosts, dtype=np.int64)
        cg = np.ascontiguousarray(col_ghosts, dtype=np.int64)
        h = C.c_void_p()
        _ck(host_lib().spmvh_matrix_create(C.byref(h), exec.h, _np_ptr(rp), _np_ptr(ci), _np_ptr(v), int(nrows_local),
                                           int(ncols_local), _np_ptr(rg), len(rg), _np_ptr(cg), len(cg),
                                           int(bool(symmetric)), int(cm)))
        return Matrix(h, exec)

    def _info(self):
        out = (_i64 * 8)()
        _ck(host_lib().spmvh_matrix_info(self.h, out))
        return list(out)

    def rows(self):
        return self._info()[0]

    def cols(self):
        return self._info()[1]

    def non_zeros(self):
        return self._info()[2]

    def symmetric(self):
        return bool(self._info()[3])

    def format_size(self):
        return self._info()[4]

    def col_map(self):
        if self._col is None:
            h = C.c_void_p()
            _ck(host_lib().spmvh_matrix_col_map(self.h, C.byref(h)))
            self._col = L2GMap(_handle=h)
        return self._col

    def row_map(self):
        if self._row is None:
            h = C.c_void_p()
            _ck(host_lib().spmvh_matrix_row_map(self.h, C.byref(h)))
            self._row = L2GMap(_handle=h)
        return self._row

    def block(self, which):
        """Host copy of a block: dict(rows, cols, rowptr, colind, values, diagonal) or None."""
        dims = (_i64 * 4)()
        _ck(host_lib().spmvh_matrix_block_dims(self.h, which, dims))
        rows, cols, nnz, has_diag = list(dims)
        if nnz < 0:
            return None
        rp = np.zeros(rows + 1, np.int32)
        ci = np.zeros(nnz, np.int32)
        v = np.zeros(nnz, np.float64)
        d = np.zeros(rows, np.float64) if has_diag else None
        _ck(host_lib().spmvh_matrix_block_copy(self.h, which, _np_ptr(rp), _np_ptr(ci), _np_ptr(v), _np_ptr(d)))
        return {"rows": rows, "cols": cols, "rowptr": rp, "colind": ci, "values": v, "diagonal": d}

    def local_block(self):
        return self.block(0)

    def remote_block(self):
        return self.block(1)

    def mult(self, x, y):
        _ck(host_lib().spmvh_matrix_mult_f64(self.h, _dev_ptr(x), _dev_ptr(y)))

    def mult_host(self, x_host, y_host):
        """Matrix::mult_host: numpy (ideally pinned) vectors in, copies pipelined with the kernel."""
        assert x_host.dtype == np.float64 and y_host.dtype == np.float64
        assert x_host.flags.c_contiguous and y_host.flags.c_contiguous
        _ck(host_lib().spmvh_matrix_mult_host_f64(self.h, _np_ptr(x_host), _np_ptr(y_host)))

    def tune(self, which=0, variant=0, tile_bytes=49152, stages=2, consumer_warps=8, ctas_per_sm=2):
        _ck(host_lib().spmvh_matrix_tune(self.h, which, variant, tile_bytes, stages, consumer_warps, ctas_per_sm))

    def close(self):
        for m in (self._col, self._row):
            if m is not None:
                m.close()
        self._col = self._row = None
        if self.h:
            host_lib().spmvh_matrix_destroy(self.h)
            self.h = None


class MatrixF32:
    """spmv::Matrix<float> (Matrix.cpp:565): create_matrix, col_map, mult, mult_host on float32 vectors."""

    def __init__(self, handle, exec):
        self.h, self.exec, self._col = handle, exec, None

    @staticmethod
    def create_matrix(exec, rowptr, colind, values, nrows_local, ncols_local, row_ghosts, col_ghosts,
                      symmetric=False, cm=COLLECTIVE_BLOCKING):
        rp = np.ascontiguousarray(rowptr, dtype=np.int32)
        ci = np.ascontiguousarray(colind, dtype=np.int32)
        v = np.ascontiguousarray(values, dtype=np.float32)
        rg = np.ascontiguousarray(row_ghosts, dtype=np.int64)
        cg = np.ascontiguousarray(col_ghosts, dtype=np.int64)
        h = C.c_void_p()
        _ck(host_lib().spmvh_matrix_create_f32(C.byref(h), exec.h, _np_ptr(rp), _np_ptr(ci), _np_ptr(v),
                                               int(nrows_local), int(ncols_local), _np_ptr(rg), len(rg), _np_ptr(cg),
                                               len(cg), int(bool(symmetric)), int(cm)))
        return MatrixF32(h, exec)

    def _info(self):
        out = (_i64 * 8)()
        _ck(host_lib().spmvh_matrix_info_f32(self.h, out))
        return list(out)

    def rows(self):
        return self._info()[0]

    def non_zeros(self):
        return self._info()[2]

    def col_map(self):
        if self._col is None:
            h = C.c_void_p()
            _ck(host_lib().spmvh_matrix_col_map_f32(self.h, C.byref(h)))
            self._col = L2GMap(_handle=h)
        return self._col

    def mult(self, x, y):
        _ck(host_lib().spmvh_matrix_mult_f32(self.h, _dev_ptr(x), _dev_ptr(y)))

    def mult_host(self, x_host, y_host):
        assert x_host.dtype == np.float32 and y_host.dtype == np.float32
        _ck(host_lib().spmvh_matrix_mult_host_f32(self.h, _np_ptr(x_host), _np_ptr(y_host)))

    def close(self):
        if self._col is not None:
            self._col.close()
            self._col = None
        if self.h:
            host_lib().spmvh_matrix_destroy_f32(self.h)
            self.h = None


def read_petsc_binary_matrix(filename, exec, symmetric=False, cm=COLLECTIVE_BLOCKING):
    h = C.c_void_p()
    _ck(host_lib().spmvh_matrix_read_petsc(C.byref(h), exec.h, os.fsencode(filename), int(bool(symmetric)), int(cm)))
    return Matrix(h, exec)


def read_petsc_binary_vector(exec, filename):
    """Returns (device pointer, local length); free with exec.free()."""
    p, n = C.c_void_p(), _i64()
    _ck(host_lib().spmvh_read_petsc_vector(exec.h, os.fsencode(filename), C.byref(p), C.byref(n)))
    return p.value, n.value


def write_petsc_binary_matrix(filename, nrows, ncols, rowptr, colind, values):
    rp = np.ascontiguousarray(rowptr, dtype=np.int32)
    ci = np.ascontiguousarray(colind, dtype=np.int32)
    v = np.ascontiguousarray(values, dtype=np.float64)
    _ck(host_lib().spmvh_write_petsc_matrix(os.fsencode(filename), nrows, ncols, _np_ptr(rp), _np_ptr(ci), _np_ptr(v)))


def write_petsc_binary_vector(filename, values):
    v = np.ascontiguousarray(values, dtype=np.float64)
    _ck(host_lib().spmvh_write_petsc_vector(os.fsencode(filename), len(v), _np_ptr(v)))


def powerlaw_rows(n, r0, r1, seed=20261104, alpha=2.0, xmin=4.0, max_len=4096, band=32768, p_global=0.1):
    """demos/CreatePowerLaw.h: rows [r0, r1) of the config-5 matrix from the host-core generator
    (same definition and defaults as synth.powerlaw_rows) -> (rowptr int64, global cols int64, values)."""
    spec = (int(n), int(seed), float(alpha), float(xmin), int(max_len), int(band), float(p_global))
    counts = np.zeros(max(r1 - r0, 1), np.int32)
    _ck(host_lib().spmvh_powerlaw_counts(*spec, int(r0), int(r1), _np_ptr(counts)))
    rowptr = np.zeros(r1 - r0 + 1, np.int64)
    np.cumsum(counts[:r1 - r0], out=rowptr[1:])
    col = np.zeros(max(int(rowptr[-1]), 1), np.int64)
    val = np.zeros(max(int(rowptr[-1]), 1), np.float64)
    _ck(host_lib().spmvh_powerlaw_fill(*spec, int(r0), int(r1), _np_ptr(rowptr), _np_ptr(col), _np_ptr(val)))
    return rowptr, col[:rowptr[-1]], val[:rowptr[-1]]


def write_powerlaw_petsc(filename, n, seed=20261104, alpha=2.0, xmin=4.0, max_len=4096, band=32768, p_global=0.1):
    """The whole n x n config-5 matrix as a PETSc binary file (input of read_petsc_binary_matrix);
    returns the entry count."""
    nnz = _i64()
    _ck(host_lib().spmvh_write_powerlaw_petsc(os.fsencode(filename), int(n), int(seed), float(alpha), float(xmin),
                                              int(max_len), int(band), float(p_global), C.byref(nnz)))
    return nnz.value


def create_A(N, exec):
    h = C.c_void_p()
    _ck(host_lib().spmvh_matrix_create_A(C.byref(h), exec.h, int(N)))
    return Matrix(h, exec)


def create_laplace3d(exec, nx, ny, nz_global, stencil, symmetric=False, cm=COLLECTIVE_BLOCKING):
    """demos/CreateLaplace3D.h: the 7-/27-point operator generated straight into executor memory
    (plane split by owner_ranges over the world's ranks); exec must be a B200Executor."""
    h = C.c_void_p()
    _ck(host_lib().spmvh_matrix_create_laplace3d(C.byref(h), exec.h, int(nx), int(ny), int(nz_global), int(stencil),
                                                 int(bool(symmetric)), int(cm)))
    return Matrix(h, exec)


def cg(exec, A, b, x, kmax, rtol, history=False):
    """spmv::cg(comm, exec, A, b, x, kmax, rtol) -> iterations [, residual history]."""
    hist = np.zeros(max(int(kmax), 1), np.float64) if history else None
    it = C.c_int()
    _ck(host_lib().spmvh_cg(exec.h, A.h, _dev_ptr(b), _dev_ptr(x), int(kmax), float(rtol), _np_ptr(hist),
                            0 if hist is None else len(hist), C.byref(it)))
    if history:
        return it.value, hist[: it.value].copy()
    return it.value


def cg_phases(exec, A, b, x, kmax, rtol=0.0):
    """spmv::cg in profile mode -> (iterations, {"update","local","remote","vector"}: device us per iteration)."""
    out = (_f64 * 4)()
    it = C.c_int()
    _ck(host_lib().spmvh_cg_phases(exec.h, A.h, _dev_ptr(b), _dev_ptr(x), int(kmax), float(rtol), out, C.byref(it)))
    return it.value, dict(zip(("update", "local", "remote", "vector"), [float(v) for v in out]))
