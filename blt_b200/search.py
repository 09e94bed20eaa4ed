"""ctypes binding of include/bltcu.h (the product's C ABI).  Never imports oracle/."""
from __future__ import annotations

import ctypes
import os

import numpy as np

BLT_STATUS_OK, BLT_STATUS_ERROR, BLT_STATUS_LP_FAIL, BLT_STATUS_INPUT_ERROR = 0, 1, 2, 3  # libblt/codes.h:5-11
BLT_CHECK_SAT, BLT_CHECK_UNSAT = 10, 11                                                    # libblt/codes.h:13-18
MAX_DIM = 96

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


class Opts(ctypes.Structure):
    _fields_ = [("device", ctypes.c_int), ("rank", ctypes.c_int), ("world", ctypes.c_int),
                ("split_depth", ctypes.c_int), ("donate_min_level", ctypes.c_int), ("no_early_exit", ctypes.c_int),
                ("warps_per_cta", ctypes.c_int), ("ctas", ctypes.c_int), ("timeout_ms", ctypes.c_int),
                ("eager", ctypes.c_int), ("sound", ctypes.c_int), ("gpus", ctypes.c_int), ("stop_flag", ctypes.c_void_p)]


class Stats(ctypes.Structure):
    _fields_ = [("niter", ctypes.c_uint64), ("nprune", ctypes.c_uint64), ("nodes", ctypes.c_uint64),
                ("pivots", ctypes.c_uint64), ("quick_prunes", ctypes.c_uint64), ("refactors", ctypes.c_uint64),
                ("cold_starts", ctypes.c_uint64), ("donated", ctypes.c_uint64),
                ("hist", ctypes.c_uint64 * (MAX_DIM + 1)), ("kernel_ms", ctypes.c_double),
                ("total_ms", ctypes.c_double), ("found", ctypes.c_int), ("launches", ctypes.c_int),
                ("warps", ctypes.c_int), ("ctas", ctypes.c_int), ("scratch_bytes", ctypes.c_uint64),
                ("busy_frac", ctypes.c_double), ("best_key", ctypes.c_uint16 * MAX_DIM), ("best_valid", ctypes.c_int),
                ("stopped", ctypes.c_int), ("gpus", ctypes.c_int), ("inband", ctypes.c_uint64)]


def lib() -> ctypes.CDLL:
    """Loads blt_b200/lib/libbltcu.so.  Fails loudly when the CUDA library has not been built."""
    global _LIB
    if _LIB is None:
        path = os.environ.get("BLTCU_LIB") or os.path.join(_HERE, "lib", "libbltcu.so")  # BLTCU_LIB: developer builds
        if not os.path.exists(path):
            raise RuntimeError(f"{path} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                               "(nvcc, sm_100a).  blt_b200 has no CPU fallback.")
        L = ctypes.CDLL(path)
        i64p, u64p, f64p = (ctypes.POINTER(t) for t in (ctypes.c_int64, ctypes.c_uint64, ctypes.c_double))
        L.bltcu_closest.argtypes = [i64p, i64p, ctypes.c_int, ctypes.c_int, i64p, u64p, u64p]
        L.bltcu_closest.restype = ctypes.c_int
        L.bltcu_closest_ex.argtypes = [i64p, i64p, ctypes.c_int, ctypes.c_int, ctypes.POINTER(Opts), i64p,
                                       ctypes.POINTER(Stats)]
        L.bltcu_closest_ex.restype = ctypes.c_int
        L.bltcu_closest_dev.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int,
                                        ctypes.POINTER(Opts), i64p, ctypes.POINTER(Stats)]
        L.bltcu_closest_dev.restype = ctypes.c_int
        L.bltcu_lll.argtypes = [i64p, i64p, ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_int)]
        L.bltcu_lll.restype = ctypes.c_int
        L.bltcu_lll_batch.argtypes = [i64p, i64p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_int),
                                      ctypes.c_int, ctypes.POINTER(Stats)]
        L.bltcu_lll_batch.restype = ctypes.c_int
        L.bltcu_lll_arbitrary.argtypes = [i64p, i64p, ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_int)]
        L.bltcu_lll_arbitrary.restype = ctypes.c_int
        L.bltcu_lll_fixed_host.argtypes = [i64p, i64p, ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_int), ctypes.c_int]
        L.bltcu_lll_fixed_host.restype = ctypes.c_int
        L.bltcu_lll_fp_host.argtypes = [i64p, i64p, ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_int)]
        L.bltcu_lll_fp_host.restype = ctypes.c_int
        L.bltcu_check_soln.argtypes = [i64p, i64p, ctypes.c_int, ctypes.c_int, i64p, ctypes.POINTER(ctypes.c_int)]
        L.bltcu_check_soln.restype = ctypes.c_int
        L.bltcu_yprune.argtypes = [i64p, ctypes.c_int, ctypes.c_int, i64p, ctypes.c_int, i64p, ctypes.POINTER(ctypes.c_int)]
        L.bltcu_yprune.restype = ctypes.c_int
        L.bltcu_infdist.argtypes = [f64p, ctypes.c_int, ctypes.c_int, i64p, ctypes.c_int, f64p, f64p, f64p]
        L.bltcu_infdist.restype = ctypes.c_int
        L.bltcu_closest_noalloc.argtypes = [i64p, i64p, ctypes.c_int, ctypes.c_int, ctypes.c_int, i64p, u64p, u64p,
                                            ctypes.POINTER(ctypes.c_int)]
        L.bltcu_closest_noalloc.restype = ctypes.c_int
        L.bltcu_closest_batch.argtypes = [i64p, i64p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.POINTER(Opts), i64p,
                                          ctypes.POINTER(ctypes.c_int), u64p, u64p, ctypes.POINTER(ctypes.c_int),
                                          ctypes.POINTER(Stats)]
        L.bltcu_closest_batch.restype = ctypes.c_int
        L.bltcu_fp64_peak.argtypes = [ctypes.c_int, ctypes.c_int, f64p]
        L.bltcu_fp64_peak.restype = ctypes.c_int
        L.bltcu_int32_peak.argtypes = [ctypes.c_int, ctypes.c_int, f64p]
        L.bltcu_int32_peak.restype = ctypes.c_int
        L.bltcu_release.argtypes = []
        L.bltcu_release.restype = None
        L.bltcu_version.argtypes = []
        L.bltcu_version.restype = ctypes.c_char_p
        _LIB = L
    return _LIB


def version() -> str:
    return lib().bltcu_version().decode()


def _i64(a):
    return np.ascontiguousarray(a, dtype=np.int64)


def _p(a, t):
    return a.ctypes.data_as(ctypes.POINTER(t))


def _stats_dict(rc, sol, st, n):
    return dict(rc=rc, sol=sol, niter=int(st.niter), nprune=int(st.nprune), nodes=int(st.nodes),
                pivots=int(st.pivots), quick=int(st.quick_prunes), refactors=int(st.refactors),
                colds=int(st.cold_starts), donated=int(st.donated),
                hist=np.array(list(st.hist)[:n + 1], dtype=np.uint64), kernel_ms=float(st.kernel_ms),
                total_ms=float(st.total_ms), found=bool(st.found), launches=int(st.launches), warps=int(st.warps),
                ctas=int(st.ctas), scratch_bytes=int(st.scratch_bytes), busy_frac=float(st.busy_frac),
                best_key=np.array(list(st.best_key)[:n], dtype=np.int64) if st.best_valid else None,
                stopped=bool(st.stopped), gpus=int(st.gpus), inband=int(st.inband))


def closest(lat, con, **opts):
    """closest() of libblt/closest.h:50-54 on the GPU, host buffers in and out.
    opts: device, rank, world, split_depth, donate_min_level, no_early_exit, warps_per_cta, ctas."""
    lat = _i64(lat)
    con = _i64(con)
    if lat.ndim != 2 or con.ndim != 2 or con.shape[1] != 2 or con.shape[0] != lat.shape[1]:
        return dict(rc=BLT_STATUS_INPUT_ERROR, sol=None, niter=0, nprune=0)  # closest.cc:173-177
    n, m = lat.shape
    if n == 0 or m == 0:
        return dict(rc=BLT_STATUS_INPUT_ERROR, sol=None, niter=0, nprune=0)
    sol = np.zeros(n, dtype=np.int64)
    o = Opts(**opts)
    st = Stats()
    rc = lib().bltcu_closest_ex(_p(lat, ctypes.c_int64), _p(con, ctypes.c_int64), n, m, ctypes.byref(o),
                                _p(sol, ctypes.c_int64), ctypes.byref(st))
    return _stats_dict(rc, sol, st, n)


def closest_dev(d_lat_ptr: int, d_con_ptr: int, n: int, m: int, **opts):
    """Same search with lat/con already in device memory (raw device pointers, e.g. torch tensor .data_ptr())."""
    sol = np.zeros(n, dtype=np.int64)
    o = Opts(**opts)
    st = Stats()
    rc = lib().bltcu_closest_dev(ctypes.c_void_p(d_lat_ptr), ctypes.c_void_p(d_con_ptr), n, m, ctypes.byref(o),
                                 _p(sol, ctypes.c_int64), ctypes.byref(st))
    return _stats_dict(rc, sol, st, n)


def closest_batch(lats, cons, **opts):
    """Batch form of closest() (bltcu_closest_batch): lats is count x n x m, cons is count x m x 2.  Every problem is
    searched on its own in the reference's depth-first order; returns per-problem arrays rc, sol, niter, nprune, found
    plus the summed statistics."""
    lats = _i64(lats)
    cons = _i64(cons)
    if lats.ndim != 3 or cons.ndim != 3 or cons.shape[2] != 2 or cons.shape[1] != lats.shape[2] or cons.shape[0] != lats.shape[0]:
        raise ValueError("closest_batch wants lats (count, n, m) and cons (count, m, 2)")
    count, n, m = lats.shape
    sol = np.zeros((count, n), dtype=np.int64)
    status = np.full(count, -1, dtype=np.int32)
    found = np.zeros(count, dtype=np.int32)
    niter = np.zeros(count, dtype=np.uint64)
    nprune = np.zeros(count, dtype=np.uint64)
    o = Opts(**opts)
    st = Stats()
    rc = lib().bltcu_closest_batch(_p(lats, ctypes.c_int64), _p(cons, ctypes.c_int64), count, n, m, ctypes.byref(o),
                                   _p(sol, ctypes.c_int64), _p(status, ctypes.c_int), _p(niter, ctypes.c_uint64),
                                   _p(nprune, ctypes.c_uint64), _p(found, ctypes.c_int), ctypes.byref(st))
    d = _stats_dict(rc, None, st, n)
    d.update(rc=rc, status=status, sol=sol, niter=niter, nprune=nprune, found=found.astype(bool))
    return d


def check_soln(lat, con, u) -> bool:
    """check_soln() of libblt/closest.cc:330-340 on the GPU."""
    lat = _i64(lat)
    con = _i64(con)
    u = _i64(u)
    n, m = lat.shape
    ok = ctypes.c_int(0)
    rc = lib().bltcu_check_soln(_p(lat, ctypes.c_int64), _p(con, ctypes.c_int64), n, m, _p(u, ctypes.c_int64),
                                ctypes.byref(ok))
    if rc != BLT_STATUS_OK:
        raise RuntimeError(f"bltcu_check_soln failed with status {rc}")
    return bool(ok.value)


def yprune(lat, u, k: int, con):
    """yprune() of libblt/yices_wrapper.h:26-34 without Yices: the exact decision whether the layer with u[k:] fixed meets
    the cube (host, Fourier-Motzkin in 128-bit integers, k <= 6).  Returns (rc, decision): rc LP_OK 20 / LP_ERROR 21,
    decision BLT_CHECK_SAT 10 (keep) / BLT_CHECK_UNSAT 11 (prune) / BLT_CHECK_UNKNOWN 12."""
    lat = _i64(lat)
    u = _i64(u)
    con = _i64(con)
    n, m = lat.shape
    d = ctypes.c_int(0)
    rc = lib().bltcu_yprune(_p(lat, ctypes.c_int64), n, m, _p(u, ctypes.c_int64), k, _p(con, ctypes.c_int64), ctypes.byref(d))
    return rc, int(d.value)


def infdist(B, u, k: int, p):
    """infdist() of libblt/glpk_wrapper.h:26-34 on the GPU (one cold solve).  Returns (rc, dist, x[0:k])."""
    B = np.ascontiguousarray(B, dtype=np.float64)
    u = _i64(u)
    p = np.ascontiguousarray(p, dtype=np.float64)
    n, m = B.shape
    x = np.zeros(max(n, 1), dtype=np.float64)
    d = ctypes.c_double(0.0)
    rc = lib().bltcu_infdist(_p(B, ctypes.c_double), n, m, _p(u, ctypes.c_int64), k, _p(p, ctypes.c_double),
                             _p(x, ctypes.c_double), ctypes.byref(d))
    return rc, float(d.value), x[:k].copy()


def fp64_peak(device: int = 0, reps: int = 5) -> float:
    """Measured FP64 FMA peak of the device in TFLOP/s (roofline denominator of the search kernel)."""
    v = ctypes.c_double(0.0)
    rc = lib().bltcu_fp64_peak(device, reps, ctypes.byref(v))
    if rc != BLT_STATUS_OK:
        raise RuntimeError(f"bltcu_fp64_peak failed with status {rc}")
    return float(v.value)


def int32_peak(device: int = 0, reps: int = 5) -> float:
    """Measured INT32 multiply-add peak of the device in Tera-operations/s (2 per IMAD)."""
    v = ctypes.c_double(0.0)
    rc = lib().bltcu_int32_peak(device, reps, ctypes.byref(v))
    if rc != BLT_STATUS_OK:
        raise RuntimeError(f"bltcu_int32_peak failed with status {rc}")
    return float(v.value)


NOALLOC_MAX_DIM = 20


def closest_noalloc(lat, con, device: int = 0):
    """closest() through bltcu_closest_noalloc: same search, but the library neither allocates nor frees device
    memory (for hosts that own the GPU heap themselves).  n, m <= 20.  Calls must not overlap on one GPU."""
    lat = _i64(lat)
    con = _i64(con)
    if lat.ndim != 2 or con.ndim != 2 or con.shape[1] != 2 or con.shape[0] != lat.shape[1]:
        return dict(rc=BLT_STATUS_INPUT_ERROR, sol=None, niter=0, nprune=0, found=False)
    n, m = lat.shape
    sol = np.zeros(max(n, 1), dtype=np.int64)
    niter, nprune, found = ctypes.c_uint64(0), ctypes.c_uint64(0), ctypes.c_int(0)
    rc = lib().bltcu_closest_noalloc(_p(lat, ctypes.c_int64), _p(con, ctypes.c_int64), n, m, device,
                                     _p(sol, ctypes.c_int64), ctypes.byref(niter), ctypes.byref(nprune),
                                     ctypes.byref(found))
    return dict(rc=rc, sol=sol[:n], niter=int(niter.value), nprune=int(nprune.value), found=bool(found.value))


def lll(B):
    """LLL() of libblt/ntl_wrapper.h:20 (exact, delta = 3/4).  Returns (rank, B_reduced, U) as int64 arrays."""
    B = _i64(B).copy()
    r, c = B.shape
    U = np.zeros((r, r), dtype=np.int64)
    rk = ctypes.c_int(0)
    rc = lib().bltcu_lll(_p(B, ctypes.c_int64), _p(U, ctypes.c_int64), r, c, ctypes.byref(rk))
    if rc != BLT_STATUS_OK:
        raise RuntimeError(f"bltcu_lll failed with status {rc}")
    return int(rk.value), B, U


def lll_batch(Bs, device: int = 0):
    """Batched LLL on the GPU (bltcu_lll_batch): Bs is count x r x c.  Returns (ranks, reduced Bs, Us, info) with the same
    bases and transforms a loop over lll() gives; info has kernel_ms, total_ms, deferred (problems the fp64-decision pass
    left to the 512-bit kernel) and on_host (problems whose numbers outgrew 512 bits: reduced by the host routine)."""
    Bs = _i64(Bs).copy()
    if Bs.ndim != 3:
        raise ValueError("lll_batch wants Bs (count, r, c)")
    count, r, c = Bs.shape
    Us = np.zeros((count, r, r), dtype=np.int64)
    ranks = np.zeros(count, dtype=np.int32)
    st = Stats()
    rc = lib().bltcu_lll_batch(_p(Bs, ctypes.c_int64), _p(Us, ctypes.c_int64), count, r, c, _p(ranks, ctypes.c_int), device,
                               ctypes.byref(st))
    if rc != BLT_STATUS_OK:
        raise RuntimeError(f"bltcu_lll_batch failed with status {rc}")
    return ranks, Bs, Us, dict(kernel_ms=float(st.kernel_ms), total_ms=float(st.total_ms), on_host=int(st.found),
                              deferred=int(st.donated), launches=int(st.launches))


def lll_fixed_host(B, limbs: int = 16):
    """Test seam: the kernel's fixed-capacity LLL executed on the host.  Returns (rc, rank, B, U)."""
    B = _i64(B).copy()
    r, c = B.shape
    U = np.zeros((r, r), dtype=np.int64)
    rk = ctypes.c_int(0)
    rc = lib().bltcu_lll_fixed_host(_p(B, ctypes.c_int64), _p(U, ctypes.c_int64), r, c, ctypes.byref(rk), limbs)
    return rc, int(rk.value), B, U


def lll_arbitrary(B):
    """Test seam: the arbitrary-precision LLL alone (lll() tries the fixed-capacity arithmetic first).  (rank, B, U)."""
    B = _i64(B).copy()
    r, c = B.shape
    U = np.zeros((r, r), dtype=np.int64)
    rk = ctypes.c_int(0)
    rc = lib().bltcu_lll_arbitrary(_p(B, ctypes.c_int64), _p(U, ctypes.c_int64), r, c, ctypes.byref(rk))
    if rc != BLT_STATUS_OK:
        raise RuntimeError(f"bltcu_lll_arbitrary failed with status {rc}")
    return int(rk.value), B, U


def lll_fp_host(B):
    """Test seam: the floating-point pass of bltcu_lll_batch (decisions in fp64 with margins, row operations in exact
    int64) executed on the host.  Returns (deferred, rank, B, U); deferred = True means the problem was left to the
    multiprecision routine and B is returned unchanged."""
    B = _i64(B).copy()
    r, c = B.shape
    U = np.zeros((r, r), dtype=np.int64)
    rk = ctypes.c_int(0)
    rc = lib().bltcu_lll_fp_host(_p(B, ctypes.c_int64), _p(U, ctypes.c_int64), r, c, ctypes.byref(rk))
    if rc not in (0, 1):
        raise RuntimeError(f"bltcu_lll_fp_host failed with status {rc}")
    return bool(rc), int(rk.value), B, U
