"""ORACLE — TEST INFRASTRUCTURE ONLY.

ctypes binding of oracle/liboracle.so (the CPU restatement of the reference algorithm) and of
oracle/_ref/libref_intervals.so (the reference's own Intervals).  Only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None
_REF = None

dp = C.POINTER(C.c_double)
ip = C.POINTER(C.c_int32)
u8p = C.POINTER(C.c_uint8)
i64p = C.POINTER(C.c_int64)
u64p = C.POINTER(C.c_uint64)


def build():
    subprocess.check_call(["make", "-s", "-C", HERE])


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(HERE, "liboracle.so")
        if not os.path.exists(path):
            build()
        _LIB = C.CDLL(path)
        _LIB.oracle_last_error.restype = C.c_char_p
        _LIB.oracle_shape_distance.restype = C.c_double
        _LIB.oracle_knn_within_ball.restype = C.c_int64
    return _LIB


def ref_intervals():
    global _REF
    if _REF is None:
        path = os.path.join(HERE, "_ref", "libref_intervals.so")
        if not os.path.exists(path):
            return None
        _REF = C.CDLL(path)
    return _REF


def _p(a, t):
    return None if a is None else a.ctypes.data_as(t)


class Oracle:
    def __init__(self, model):
        self.model = model
        self._desc = model.struct()
        self.h = C.c_void_p()
        if lib().oracle_create(C.byref(self._desc), C.byref(self.h)) != 0:
            raise RuntimeError(lib().oracle_last_error().decode())
        self.nq = lib().oracle_nq(self.h)
        self.nv = lib().oracle_nv(self.h)

    def __del__(self):
        if getattr(self, "h", None):
            try:
                lib().oracle_destroy(self.h)
            except TypeError:  # interpreter shutdown: module globals are already gone
                pass
            self.h = None

    # tables
    def config_pairs(self):
        n = lib().oracle_num_config_pairs(self.h)
        a, b, m = np.zeros(n, np.int32), np.zeros(n, np.int32), np.zeros(n)
        lib().oracle_get_config_pairs(self.h, _p(a, ip), _p(b, ip), _p(m, dp))
        return a, b, m

    def body_pairs(self):
        n = lib().oracle_num_body_pairs(self.h)
        ja, jb = np.zeros(n, np.int32), np.zeros(n, np.int32)
        pb, cb, rb = (np.zeros(n + 1, np.int32) for _ in range(3))
        lib().oracle_get_body_pairs(self.h, _p(ja, ip), _p(jb, ip), _p(pb, ip), _p(cb, ip), _p(rb, ip))
        ng = lib().oracle_num_body_geom_pairs(self.h)
        ga, gb, gm = np.zeros(ng, np.int32), np.zeros(ng, np.int32), np.zeros(ng)
        lib().oracle_get_body_geom_pairs(self.h, _p(ga, ip), _p(gb, ip), _p(gm, dp))
        out = dict(joint_a=ja, joint_b=jb, pair_begin=pb, coef_begin=cb, rcoef_begin=rb,
                   geomA=ga, geomB=gb, margin=gm)
        for rev, key in ((0, "coef"), (1, "rcoef")):
            nc = lib().oracle_num_coefficients(self.h, rev)
            j, v = np.zeros(nc, np.int32), np.zeros(nc)
            lib().oracle_get_coefficients(self.h, rev, _p(j, ip), _p(v, dp))
            out[key + "_joint"], out[key + "_value"] = j, v
        return out

    def body_radii(self):
        r = np.zeros(self.model.njoints + 1)
        lib().oracle_get_body_radii(self.h, _p(r, dp))
        return r

    # FK / distances
    def geometry_placements(self, q):
        q = np.ascontiguousarray(q, np.float64).reshape(-1, self.nq)
        out = np.zeros((q.shape[0], self.model.ngeoms, 12))
        lib().oracle_geometry_placements(self.h, _p(q, dp), C.c_size_t(q.shape[0]), _p(out, dp))
        return out

    def pair_distances(self, q):
        q = np.ascontiguousarray(q, np.float64).reshape(-1, self.nq)
        npairs = lib().oracle_num_config_pairs(self.h)
        out = np.zeros((q.shape[0], npairs))
        lib().oracle_pair_distances(self.h, _p(q, dp), C.c_size_t(q.shape[0]), _p(out, dp))
        return out

    def all_collisions(self, q):
        q = np.ascontiguousarray(q, np.float64).reshape(-1, self.nq)
        n = q.shape[0]
        npairs = lib().oracle_num_config_pairs(self.h)
        words = (npairs + 31) // 32
        valid, first, count = np.zeros(n, np.uint8), np.zeros(n, np.int32), np.zeros(n, np.int32)
        mask = np.zeros((n, words), np.uint32)
        lib().oracle_all_collisions(self.h, _p(q, dp), C.c_size_t(n), _p(valid, u8p), _p(first, ip), mask.ctypes.data_as(C.c_void_p), _p(count, ip))
        bits = np.unpackbits(mask.view(np.uint8), axis=1, bitorder="little")[:, :npairs]
        return valid, first, [np.nonzero(b)[0] for b in bits], count

    def distance_between_objects(self, q):
        q = np.ascontiguousarray(q, np.float64).reshape(-1, self.nq)
        per = lib().oracle_num_distance_pairs(self.h)
        out = np.zeros((q.shape[0], per))
        lib().oracle_distance_between_objects(self.h, _p(q, dp), C.c_size_t(q.shape[0]), _p(out, dp))
        return out

    def validate_configs(self, q, check_bounds=False, nthreads=0):
        q = np.ascontiguousarray(q, np.float64).reshape(-1, self.nq)
        n = q.shape[0]
        valid, pair, reason = np.zeros(n, np.uint8), np.zeros(n, np.int32), np.zeros(n, np.uint8)
        lib().oracle_validate_configs(self.h, _p(q, dp), C.c_size_t(n), C.c_uint32(1 if check_bounds else 0),
                                      _p(valid, u8p), _p(pair, ip), _p(reason, u8p), C.c_int(nthreads))
        return valid, pair, reason

    def straight_lengths(self, q1, q2):
        q1 = np.ascontiguousarray(q1, np.float64).reshape(-1, self.nq)
        q2 = np.ascontiguousarray(q2, np.float64).reshape(-1, self.nq)
        L = np.zeros(q1.shape[0])
        lib().oracle_straight_lengths(self.h, _p(q1, dp), _p(q2, dp), C.c_size_t(q1.shape[0]), _p(L, dp))
        return L

    def interpolate(self, q1, q2, L, t):
        q1 = np.ascontiguousarray(q1, np.float64).reshape(-1, self.nq)
        q2 = np.ascontiguousarray(q2, np.float64).reshape(-1, self.nq)
        L, t = np.ascontiguousarray(L, np.float64), np.ascontiguousarray(t, np.float64)
        q = np.zeros_like(q1)
        lib().oracle_interpolate(self.h, _p(q1, dp), _p(q2, dp), _p(L, dp), _p(t, dp), C.c_size_t(q1.shape[0]), _p(q, dp))
        return q

    def difference(self, q1, q2):
        """q2 (-) q1, [n][nv]"""
        q1 = np.ascontiguousarray(q1, np.float64).reshape(-1, self.nq)
        q2 = np.ascontiguousarray(q2, np.float64).reshape(-1, self.nq)
        v = np.zeros((q1.shape[0], self.nv))
        lib().oracle_difference(self.h, _p(q1, dp), _p(q2, dp), C.c_size_t(q1.shape[0]), _p(v, dp))
        return v

    def validate_edges(self, method, param, q1, q2, L, reverse=False, check_bounds=False, order_mode=0,
                       max_iter=0, nthreads=0):
        q1 = np.ascontiguousarray(q1, np.float64).reshape(-1, self.nq)
        q2 = np.ascontiguousarray(q2, np.float64).reshape(-1, self.nq)
        L = np.ascontiguousarray(L, np.float64)
        n = q1.shape[0]
        valid, reason = np.zeros(n, np.uint8), np.zeros(n, np.uint8)
        t_fail, t_last = np.zeros(n), np.zeros(n)
        pair = np.zeros(n, np.int32)
        samples = np.zeros(n, np.uint64)
        rc = lib().oracle_validate_edges(self.h, C.c_int(method), C.c_double(param), C.c_int(1 if reverse else 0),
                                         _p(q1, dp), _p(q2, dp), _p(L, dp), C.c_size_t(n),
                                         C.c_uint32(1 if check_bounds else 0), _p(valid, u8p), _p(t_fail, dp),
                                         _p(t_last, dp), _p(pair, ip), _p(reason, u8p), _p(samples, u64p),
                                         C.c_int(order_mode), C.c_uint64(max_iter), C.c_int(nthreads))
        if rc != 0:
            raise RuntimeError(lib().oracle_last_error().decode())
        return dict(valid=valid, t_fail=t_fail, t_last_valid=t_last, pair=pair, reason=reason, samples=samples)

    def _spline_args(self, base, params, L):
        base = np.ascontiguousarray(base, np.float64).reshape(-1, self.nq)
        params = np.ascontiguousarray(params, np.float64).reshape(base.shape[0], 4, self.nq)
        return base, params, np.ascontiguousarray(L, np.float64)

    def validate_spline_edges(self, method, param, base, params, L, reverse=False, check_bounds=False, max_iter=0, nthreads=0):
        """edges that are path::Spline<BernsteinBasis, 3>: base [n][nq], params [n][4][nv], L [n]"""
        base, params, L = self._spline_args(base, params, L)
        n = base.shape[0]
        valid, reason = np.zeros(n, np.uint8), np.zeros(n, np.uint8)
        t_fail, t_last = np.zeros(n), np.zeros(n)
        pair = np.zeros(n, np.int32)
        samples = np.zeros(n, np.uint64)
        rc = lib().oracle_validate_spline_edges(self.h, C.c_int(method), C.c_double(param), C.c_int(1 if reverse else 0),
                                                _p(base, dp), _p(params, dp), _p(L, dp), C.c_size_t(n),
                                                C.c_uint32(1 if check_bounds else 0), _p(valid, u8p), _p(t_fail, dp),
                                                _p(t_last, dp), _p(pair, ip), _p(reason, u8p), _p(samples, u64p),
                                                C.c_uint64(max_iter), C.c_int(nthreads))
        if rc != 0:
            raise RuntimeError(lib().oracle_last_error().decode())
        return dict(valid=valid, t_fail=t_fail, t_last_valid=t_last, pair=pair, reason=reason, samples=samples)

    def spline_eval(self, base, params, L, t):
        base, params, L = self._spline_args(base, params, L)
        t = np.ascontiguousarray(t, np.float64)
        q = np.zeros_like(base)
        if lib().oracle_spline_eval(self.h, _p(base, dp), _p(params, dp), _p(L, dp), _p(t, dp), C.c_size_t(base.shape[0]), _p(q, dp)) != 0:
            raise RuntimeError(lib().oracle_last_error().decode())
        return q

    def spline_velocity_bound(self, base, params, L, ta, tb):
        base, params, L = self._spline_args(base, params, L)
        ta, tb = np.ascontiguousarray(ta, np.float64), np.ascontiguousarray(tb, np.float64)
        vb = np.zeros_like(base)
        if lib().oracle_spline_velocity_bound(self.h, _p(base, dp), _p(params, dp), _p(L, dp), _p(ta, dp), _p(tb, dp),
                                              C.c_size_t(base.shape[0]), _p(vb, dp)) != 0:
            raise RuntimeError(lib().oracle_last_error().decode())
        return vb

    def validate_edge_chains(self, method, param, q1, q2, L, chain_begin, reverse=False, nthreads=0):
        q1 = np.ascontiguousarray(q1, np.float64).reshape(-1, self.nq)
        q2 = np.ascontiguousarray(q2, np.float64).reshape(-1, self.nq)
        L = np.ascontiguousarray(L, np.float64)
        cb = np.ascontiguousarray(chain_begin, np.int32)
        nc = cb.shape[0] - 1
        valid, reason = np.zeros(nc, np.uint8), np.zeros(nc, np.uint8)
        t_fail, t_last = np.zeros(nc), np.zeros(nc)
        pair, bad = np.zeros(nc, np.int32), np.zeros(nc, np.int32)
        rc = lib().oracle_validate_edge_chains(self.h, C.c_int(method), C.c_double(param), C.c_int(1 if reverse else 0),
                                               _p(q1, dp), _p(q2, dp), _p(L, dp), _p(cb, ip), C.c_size_t(nc), _p(valid, u8p),
                                               _p(t_fail, dp), _p(t_last, dp), _p(pair, ip), _p(reason, u8p), _p(bad, ip),
                                               C.c_int(nthreads))
        if rc != 0:
            raise RuntimeError(lib().oracle_last_error().decode())
        return dict(valid=valid, t_fail=t_fail, t_last_valid=t_last, pair=pair, reason=reason, first_bad_edge=bad)

    def counters(self):
        c = np.zeros(5, np.uint64)
        lib().oracle_get_counters(self.h, _p(c, u64p))
        return dict(gjk_calls=int(c[0]), gjk_iterations=int(c[1]), pair_evals=int(c[2]), samples=int(c[3]),
                    closed_form_calls=int(c[4]))

    # nearest neighbour
    def knn_append(self, q, cc=None):
        q = np.ascontiguousarray(q, np.float64).reshape(-1, self.nq)
        cc = None if cc is None else np.ascontiguousarray(cc, np.int32)
        lib().oracle_knn_append(self.h, _p(q, dp), _p(cc, ip), C.c_size_t(q.shape[0]))

    def knn_relabel(self, a, b):
        lib().oracle_knn_relabel(self.h, C.c_int(a), C.c_int(b))

    def knn_query(self, queries, k, cc=-1, nthreads=0):
        queries = np.ascontiguousarray(queries, np.float64).reshape(-1, self.nq)
        n = queries.shape[0]
        idx, dist = np.zeros((n, k), np.int64), np.zeros((n, k))
        lib().oracle_knn_query(self.h, _p(queries, dp), C.c_size_t(n), C.c_int(k), C.c_int(cc), _p(idx, i64p),
                               _p(dist, dp), C.c_int(nthreads))
        return idx, dist

    def knn_within_ball(self, q, radius, cc=-1, cap=1 << 20):
        q = np.ascontiguousarray(q, np.float64).reshape(self.nq)
        idx = np.zeros(cap, np.int64)
        n = lib().oracle_knn_within_ball(self.h, _p(q, dp), C.c_double(radius), C.c_int(cc), _p(idx, i64p),
                                         C.c_size_t(cap))
        return idx[:min(n, cap)].copy()


def sincos(x):
    x = np.ascontiguousarray(x, np.float64)
    s, c = np.zeros_like(x), np.zeros_like(x)
    lib().oracle_sincos(_p(x, dp), C.c_size_t(x.size), _p(s, dp), _p(c, dp))
    return s, c


def atan(x):
    x = np.ascontiguousarray(x, np.float64)
    y = np.zeros_like(x)
    lib().oracle_atan(_p(x, dp), C.c_size_t(x.size), _p(y, dp))
    return y


def shape_distance(ta, pa, Ta, tb, pb, Tb, va=None, vb=None):
    pa = np.ascontiguousarray(list(pa) + [0.0] * (4 - len(pa)), np.float64)
    pb = np.ascontiguousarray(list(pb) + [0.0] * (4 - len(pb)), np.float64)
    Ta, Tb = np.ascontiguousarray(Ta, np.float64), np.ascontiguousarray(Tb, np.float64)
    va = None if va is None else np.ascontiguousarray(va, np.float64)
    vb = None if vb is None else np.ascontiguousarray(vb, np.float64)
    return lib().oracle_shape_distance(C.c_int(ta), _p(pa, dp), _p(Ta, dp), _p(va, dp),
                                       C.c_int(0 if va is None else va.shape[0]), C.c_int(tb), _p(pb, dp),
                                       _p(Tb, dp), _p(vb, dp), C.c_int(0 if vb is None else vb.shape[0]))


def intervals_union(lo, hi, which="oracle", cap=4096):
    lo, hi = np.ascontiguousarray(lo, np.float64), np.ascontiguousarray(hi, np.float64)
    olo, ohi = np.zeros(cap), np.zeros(cap)
    if which == "oracle":
        n = lib().oracle_intervals_union(_p(lo, dp), _p(hi, dp), C.c_size_t(lo.size), _p(olo, dp), _p(ohi, dp), cap)
    else:
        n = ref_intervals().ref_intervals_union(_p(lo, dp), _p(hi, dp), C.c_size_t(lo.size), _p(olo, dp),
                                                _p(ohi, dp), cap)
    return np.stack([olo[:n], ohi[:n]], axis=1)
