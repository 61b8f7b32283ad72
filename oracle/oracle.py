"""ctypes wrapper of oracle/liboracle.so -- TEST INFRASTRUCTURE ONLY.

May be imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs; never by toyslam_b200 (the product).  See oracle/oracle.cpp header.
"""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build():
    subprocess.check_call(["make", "-C", _HERE, "liboracle.so"], stdout=subprocess.DEVNULL)


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(path):
            build()
        L = C.CDLL(path)
        L.orc_load_wire.restype = C.c_void_p
        L.orc_load_wire.argtypes = [C.c_char_p, C.c_uint32]
        L.orc_error.restype = C.c_char_p
        L.orc_error.argtypes = [C.c_void_p]
        L.orc_assemble_dense.restype = C.c_double
        L.orc_linearize_sparse.restype = C.c_double
        L.orc_reply_size.restype = C.c_uint32
        for f in ("orc_free", "orc_counts", "orc_ids", "orc_edge_index", "orc_get_state", "orc_set_state",
                  "orc_edge_odom", "orc_edge_lm", "orc_assemble_dense", "orc_optimize_dense",
                  "orc_optimize_sparse", "orc_linearize_sparse", "orc_reply_size", "orc_write_reply"):
            getattr(L, f).argtypes = None
        _LIB = L
    return _LIB


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class OracleGraph:
    def __init__(self, payload: bytes):
        self.L = lib()
        self.h = C.c_void_p(self.L.orc_load_wire(payload, len(payload)))
        err = self.L.orc_error(self.h).decode()
        if err:
            self.L.orc_free(self.h); self.h = None
            raise ValueError(err)
        c = np.zeros(6, np.uint32)
        self.L.orc_counts(self.h, _p(c))
        self.P, self.Lm, self.Eo, self.El, self.F, self.dim = (int(v) for v in c)

    def __del__(self):
        if getattr(self, "h", None):
            self.L.orc_free(self.h); self.h = None

    def ids(self):
        a, b = np.zeros(self.P, np.uint32), np.zeros(self.Lm, np.uint32)
        self.L.orc_ids(self.h, _p(a), _p(b)); return a, b

    def edge_index(self):
        a, b = np.zeros(self.Eo, np.uint32), np.zeros(self.Eo, np.uint32)
        c, d = np.zeros(self.El, np.uint32), np.zeros(self.El, np.uint32)
        self.L.orc_edge_index(self.h, _p(a), _p(b), _p(c), _p(d)); return a, b, c, d

    def state(self):
        a, b = np.zeros((self.P, 3)), np.zeros((self.Lm, 2))
        self.L.orc_get_state(self.h, _p(a), _p(b)); return a, b

    def set_state(self, pose_xyt, lm_xy):
        a, b = np.ascontiguousarray(pose_xyt, np.float64), np.ascontiguousarray(lm_xy, np.float64)
        self.L.orc_set_state(self.h, _p(a), _p(b))

    def edge_odom(self, k):
        e = np.zeros(3); self.L.orc_edge_odom(self.h, C.c_uint32(k), _p(e)); return e

    def edge_lm(self, k):
        e, A, B = np.zeros(2), np.zeros((2, 3)), np.zeros((2, 2))
        self.L.orc_edge_lm(self.h, C.c_uint32(k), _p(e), _p(A), _p(B)); return e, A, B

    def assemble_dense(self):
        H, b = np.zeros((self.dim, self.dim)), np.zeros(self.dim)
        err = self.L.orc_assemble_dense(self.h, _p(H), _p(b)); return H, b, float(err)

    def optimize_dense(self, iterations, enable_exits=True):
        traj = np.zeros(max(iterations, 1)); code = C.c_int(0)
        n = self.L.orc_optimize_dense(self.h, C.c_int(iterations), C.c_int(int(enable_exits)), _p(traj), C.byref(code))
        return traj[:n].copy(), code.value

    def optimize_sparse(self, iterations, enable_exits=True, tol=1e-12, max_pcg=5000, threads=0):
        traj = np.zeros(max(iterations, 1)); its = np.zeros(max(iterations, 1), np.int32); code = C.c_int(0)
        n = self.L.orc_optimize_sparse(self.h, C.c_int(iterations), C.c_int(int(enable_exits)), C.c_double(tol),
                                       C.c_int(max_pcg), C.c_int(threads), _p(traj), _p(its), C.byref(code))
        return traj[:n].copy(), its[:n].copy(), code.value

    def residual_sparse(self, delta_pose, delta_lm, threads=0):
        """(||H delta - b|| / ||b||, chi2) of the full system linearised at the oracle's current state"""
        dp = np.ascontiguousarray(delta_pose, np.float64); dl = np.ascontiguousarray(delta_lm, np.float64)
        chi2 = C.c_double(0)
        self.L.orc_residual_sparse.restype = C.c_double
        r = self.L.orc_residual_sparse(self.h, _p(dp), _p(dl), C.c_int(threads), C.byref(chi2))
        return float(r), chi2.value

    def linearize_sparse(self, reps=1, threads=0):
        return float(self.L.orc_linearize_sparse(self.h, C.c_int(reps), C.c_int(threads)))

    def reply(self) -> bytes:
        n = self.L.orc_reply_size(self.h)
        out = C.create_string_buffer(n)
        self.L.orc_write_reply(self.h, out); return out.raw


def solve_dense(H, b):
    H = np.ascontiguousarray(H, np.float64); x = np.array(b, np.float64)
    lib().orc_solve_dense(_p(H), _p(x), C.c_uint32(len(x))); return x


def max_threads():
    return int(lib().orc_max_threads())
