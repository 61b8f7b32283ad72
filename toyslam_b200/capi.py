"""ctypes binding of libtoyslam_b200.so (include/toyslam_b200.h) and a thin host-side mirror of
the reference's optimizer interface for this path:

    reference (C++)                                     here
    CreateGraph/CreateOptimizer (GraphManager.h:73-122)  GraphOptimizerGpu(iterations, ...)
    DeserializeGraphStart(graph, bytes)  (:35-54)        .load_wire(payload)
    optimizer->Optimize(graph)  (ConnectionHandler.h:23) .optimize()
    SerializeGraphStart(graph)  (:56-71)                 .reply()

The library is the product; this file only marshals arguments.  There is no CPU fallback: if the
shared library is missing or no CUDA device is present the calls raise.
"""
import ctypes as C
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libtoyslam_b200.so")

SYMBOLS = ["tsb_nccl_unique_id", "tsb_create", "tsb_destroy", "tsb_last_error", "tsb_load_wire", "tsb_load_soa", "tsb_get_counts", "tsb_host_plan", "tsb_host_structure",
           "tsb_reset_state", "tsb_set_state", "tsb_linearize", "tsb_solve", "tsb_optimize", "tsb_read_ids",
           "tsb_read_state", "tsb_read_delta", "tsb_get_transfer_bytes", "tsb_read_system", "tsb_read_structure", "tsb_reply_size",
           "tsb_write_reply", "tsb_odom_linearize_noalloc"]


class Config(C.Structure):
    _fields_ = [("device", C.c_int32), ("rank", C.c_int32), ("world", C.c_int32), ("nccl_id", C.c_void_p),
                ("pcg_tol", C.c_double), ("pcg_max_iters", C.c_int32), ("enable_exits", C.c_int32), ("verbose", C.c_int32)]


class GraphSoa(C.Structure):
    _fields_ = [("n_pose", C.c_uint32), ("n_lm", C.c_uint32), ("n_odom", C.c_uint32), ("n_lmedge", C.c_uint32),
                ("n_fixed", C.c_uint32),
                ("pose_id", C.c_void_p), ("pose_xyt", C.c_void_p), ("lm_id", C.c_void_p), ("lm_xy", C.c_void_p),
                ("odom_ij", C.c_void_p), ("odom_Z", C.c_void_p), ("odom_w", C.c_void_p),
                ("lmedge_ij", C.c_void_p), ("lmedge_meas", C.c_void_p), ("lmedge_w", C.c_void_p), ("fixed_id", C.c_void_p)]


class Counts(C.Structure):
    _fields_ = [(n, C.c_uint32) for n in ("n_pose", "n_lm", "n_odom", "n_lmedge", "n_fixed", "dim", "pose_begin",
                                          "pose_end", "n_lmedge_local", "n_odom_local", "n_chunks", "n_blocks", "n_slots",
                                          "pcg_path", "replicated", "n_aggregates", "n_ml_levels", "n_ml_top", "structure_reuses")]


class IterStats(C.Structure):
    _fields_ = [("chi2", C.c_double), ("delta_norm", C.c_double), ("pcg_iters", C.c_int32),
                ("pcg_rel_residual", C.c_double), ("ms_linearize", C.c_float), ("ms_solve", C.c_float)]


class Report(C.Structure):
    _fields_ = [("iterations_run", C.c_int32), ("exit_code", C.c_int32), ("final_chi2", C.c_double),
                ("pcg_iters_total", C.c_int64), ("kernel_launches", C.c_int64), ("ms_total", C.c_float),
                ("ms_linearize_total", C.c_float), ("ms_solve_total", C.c_float), ("ms_pcg_total", C.c_float),
                ("coarse_refreshes", C.c_int32), ("pcg_not_converged", C.c_int32), ("coarse_failures", C.c_int32),
                ("singular_blocks", C.c_int32)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        try:                      # load torch first when present so that one libnccl.so.2 serves both
            import torch  # noqa: F401
        except Exception:
            pass
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing: build it with `make -C toyslam_b200/csrc` "
                               "(or __graft_entry__.build()); there is no fallback path")
        L = C.CDLL(LIB_PATH)
        L.tsb_last_error.restype = C.c_char_p
        L.tsb_last_error.argtypes = [C.c_void_p]
        L.tsb_destroy.restype = None
        _lib = L
    return _lib


def nccl_unique_id() -> bytes:
    buf = C.create_string_buffer(128)
    rc = lib().tsb_nccl_unique_id(buf)
    if rc:
        raise TsbError(rc, lib().tsb_last_error(None).decode())
    return buf.raw


NOALLOC_MAX_ODOM_EDGES = 1048576


def odom_linearize_noalloc(device, stream, n_pose, n_edge, pose_xyt, edge_ij, edge_zinv, edge_w, hpp_off, w_e, chi2):
    """tsb_odom_linearize_noalloc: every array argument is a DEVICE pointer (int); nothing is allocated.  Raises TsbError."""
    err = C.create_string_buffer(256)
    vp = lambda a: C.c_void_p(int(a) if a else None)
    rc = lib().tsb_odom_linearize_noalloc(C.c_int32(device), vp(stream), C.c_uint32(n_pose), C.c_uint32(n_edge), vp(pose_xyt), vp(edge_ij),
                                          vp(edge_zinv), vp(edge_w), vp(hpp_off), vp(w_e), vp(chi2), err, C.c_uint32(256))
    if rc:
        raise TsbError(rc, err.value.decode())


def host_plan(payload: bytes, rank: int, world: int):
    """CPU-only: counts of rank `rank`'s share, ghost / boundary pose counts and the rank-independent hash."""
    c, g, b, h = Counts(), C.c_uint32(), C.c_uint32(), C.c_uint64()
    rc = lib().tsb_host_plan(payload, C.c_uint32(len(payload)), C.c_int32(rank), C.c_int32(world), C.byref(c),
                             C.byref(g), C.byref(b), C.byref(h))
    if rc:
        raise TsbError(rc, lib().tsb_last_error(None).decode())
    return c, g.value, b.value, h.value


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def host_structure(payload: bytes, rank: int, world: int):
    """Integer structure arrays of one rank, built on the host (no GPU): dict like GraphOptimizerGpu.structure()."""
    c, _, _, _ = host_plan(payload, rank, world)
    nP = c.pose_end - c.pose_begin
    out = dict(lm_perm=np.zeros(c.n_lmedge_local, np.uint32), pose_ptr=np.zeros(nP + 1, np.uint32),
               chunk_desc=np.zeros((c.n_chunks, 8), np.uint32), slot_lm=np.zeros(c.n_slots, np.uint32),
               lm_packed=np.zeros(c.n_lmedge_local, np.uint32))
    rc = lib().tsb_host_structure(payload, C.c_uint32(len(payload)), C.c_int32(rank), C.c_int32(world),
                                  *(_p(out[k]) for k in ("lm_perm", "pose_ptr", "chunk_desc", "slot_lm", "lm_packed")))
    if rc:
        raise TsbError(rc, lib().tsb_last_error(None).decode())
    return c, out


class TsbError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"tsb error {code}: {msg}")
        self.code = code


class GraphOptimizerGpu:
    """One context on one GPU (one process per GPU).  Mirrors OptimizerGpu<T> + GraphGpu<T>."""

    def __init__(self, iterations=10, device=0, rank=0, world=1, nccl_id=None, pcg_tol=1e-12, pcg_max_iters=4000,
                 enable_exits=True, verbose=False):
        self.L = lib()
        self.iterations = iterations
        self._id = nccl_id
        self._idbuf = C.create_string_buffer(nccl_id, 128) if nccl_id else None
        cfg = Config(device, rank, world, C.cast(self._idbuf, C.c_void_p) if nccl_id else None,
                     pcg_tol, pcg_max_iters, int(enable_exits), int(verbose))
        h = C.c_void_p()
        rc = self.L.tsb_create(C.byref(cfg), C.byref(h))
        if rc:
            raise TsbError(rc, self.L.tsb_last_error(None).decode())
        self.h = h
        self.counts = None

    def close(self):
        if getattr(self, "h", None):
            self.L.tsb_destroy(self.h); self.h = None

    __del__ = close

    def _ck(self, rc):
        if rc:
            raise TsbError(rc, self.L.tsb_last_error(self.h).decode())

    def load_wire(self, payload):
        buf = (C.c_char * len(payload)).from_buffer_copy(payload) if isinstance(payload, (bytes, bytearray)) else payload
        self._ck(self.L.tsb_load_wire(self.h, buf, C.c_uint32(len(payload))))
        c = Counts(); self._ck(self.L.tsb_get_counts(self.h, C.byref(c))); self.counts = c
        return c

    def refresh_counts(self):
        """counts again (pcg_path reports the path the LAST solve took)"""
        c = Counts(); self._ck(self.L.tsb_get_counts(self.h, C.byref(c))); self.counts = c
        return c

    def load_wire_ptr(self, ptr, nbytes):
        """payload already in (pinned) host memory at address `ptr`"""
        self._ck(self.L.tsb_load_wire(self.h, C.c_void_p(ptr), C.c_uint32(nbytes)))
        c = Counts(); self._ck(self.L.tsb_get_counts(self.h, C.byref(c))); self.counts = c
        return c

    def load_soa(self, g):
        keep = {k: np.ascontiguousarray(v) for k, v in g.items()}
        s = GraphSoa(len(keep["pose_id"]), len(keep["lm_id"]), len(keep["od_ij"]), len(keep["lm_ij"]), len(keep["fixed_ids"]),
                     *(keep[k].ctypes.data for k in ("pose_id", "pose_xyt", "lm_id", "lm_xy", "od_ij", "od_Z", "od_w",
                                                     "lm_ij", "lm_meas", "lm_w", "fixed_ids")))
        self._ck(self.L.tsb_load_soa(self.h, C.byref(s)))
        c = Counts(); self._ck(self.L.tsb_get_counts(self.h, C.byref(c))); self.counts = c
        return c

    def reset_state(self):
        self._ck(self.L.tsb_reset_state(self.h))

    def set_state(self, pose_xyt, lm_xy):
        a, b = np.ascontiguousarray(pose_xyt, np.float64), np.ascontiguousarray(lm_xy, np.float64)
        self._ck(self.L.tsb_set_state(self.h, _p(a), _p(b)))

    def linearize(self):
        st = IterStats(); self._ck(self.L.tsb_linearize(self.h, C.byref(st))); return st

    def solve(self):
        st = IterStats(); self._ck(self.L.tsb_solve(self.h, C.byref(st))); return st

    def optimize(self, iterations=None):
        n = self.iterations if iterations is None else iterations
        rep = Report(); per = (IterStats * max(n, 1))()
        self._ck(self.L.tsb_optimize(self.h, C.c_uint32(n), C.byref(rep), per))
        return rep, list(per)[:rep.iterations_run]

    def ids(self):
        a, b = np.zeros(self.counts.n_pose, np.uint32), np.zeros(self.counts.n_lm, np.uint32)
        self._ck(self.L.tsb_read_ids(self.h, _p(a), _p(b))); return a, b

    def state(self):
        a, b = np.zeros((self.counts.n_pose, 3)), np.zeros((self.counts.n_lm, 2))
        self._ck(self.L.tsb_read_state(self.h, _p(a), _p(b))); return a, b

    def delta(self):
        c = self.counts
        a, b = np.zeros((c.pose_end - c.pose_begin, 3)), np.zeros((c.n_lm, 2))
        self._ck(self.L.tsb_read_delta(self.h, _p(a), _p(b))); return a, b

    def system(self):
        c = self.counts
        out = dict(hpp_diag=np.zeros((c.n_pose, 6)), b_pose=np.zeros((c.n_pose, 3)), hpl=np.zeros((c.n_lmedge, 3, 2)),
                   hpp_off=np.zeros((c.n_odom, 3)), hll=np.zeros((c.n_lm, 3)), b_lm=np.zeros((c.n_lm, 2)))
        self._ck(self.L.tsb_read_system(self.h, *(_p(out[k]) for k in ("hpp_diag", "b_pose", "hpl", "hpp_off", "hll", "b_lm"))))
        return out

    def structure(self):
        c = self.counts
        nP = c.pose_end - c.pose_begin
        out = dict(lm_perm=np.zeros(c.n_lmedge_local, np.uint32), lm_packed=np.zeros(c.n_lmedge_local, np.uint32),
                   chunk_desc=np.zeros((c.n_chunks, 8), np.uint32), slot_lm=np.zeros(c.n_slots, np.uint32),
                   pose_ptr=np.zeros(nP + 1, np.uint32), lm_part_ptr=np.zeros(c.n_lm + 1, np.uint32),
                   lm_part_idx=np.zeros(c.n_slots, np.uint32))
        self._ck(self.L.tsb_read_structure(self.h, *(_p(out[k]) for k in ("lm_perm", "lm_packed", "chunk_desc", "slot_lm",
                                                                            "pose_ptr", "lm_part_ptr", "lm_part_idx"))))
        return out

    def transfer_bytes(self):
        a, b = C.c_uint64(), C.c_uint64()
        self._ck(self.L.tsb_get_transfer_bytes(self.h, C.byref(a), C.byref(b))); return a.value, b.value

    def write_reply_into(self, ptr, cap):
        self._ck(self.L.tsb_write_reply(self.h, C.c_void_p(ptr), C.c_uint32(cap)))

    def reply_size(self):
        n = C.c_uint32(); self._ck(self.L.tsb_reply_size(self.h, C.byref(n))); return n.value

    def reply(self) -> bytes:
        n = C.c_uint32(); self._ck(self.L.tsb_reply_size(self.h, C.byref(n)))
        out = C.create_string_buffer(n.value)
        self._ck(self.L.tsb_write_reply(self.h, out, n)); return out.raw
