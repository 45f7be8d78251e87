"""hpp-core_b200 — B200-native batched validation engine for hpp-core's hot path.

Thin ctypes binding of the C ABI (include/hppb200.h, built into hpp-core_b200/libhppb200.so)
plus a Python mirror of the reference's operator interface for this path (same names and
argument meaning as hpp::core::{CollisionValidation, JointBoundValidation, ConfigValidations,
pathValidation::Discretized, continuousValidation::{Progressive, Dichotomy},
nearestNeighbor::Basic}), extended with batch entry points.

The directory name contains a hyphen, so import it through `load()` in __graft_entry__.py
(importlib) under the module name `hpp_core_b200`.

There is NO CPU fallback: every compute call goes to the CUDA library and raises if the
library or a GPU is missing.
"""
import ctypes as C
import os
import subprocess

import numpy as np

from .model_io import (ModelDescription, freeflyer_ranks, hppb200_model_desc, normalize_quaternions,  # noqa: F401
                       uniform_configs)

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("HPPB200_LIB") or os.path.join(HERE, "libhppb200.so")  # (HPPB200_LIB: A/B builds of the same library)

OK, ERR_INVALID_ARGUMENT, ERR_RUNTIME, ERR_CUDA, ERR_UNSUPPORTED = range(5)
METHOD_DISCRETIZED, METHOD_PROGRESSIVE, METHOD_DICHOTOMY = 0, 1, 2
REASON_VALID, REASON_COLLISION, REASON_JOINT_BOUND, REASON_ITER_CAP = 0, 1, 2, 3
CHECK_JOINT_BOUNDS, DEVICE_POINTERS = 1, 2

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int32)
_u8p = C.POINTER(C.c_uint8)
_i64p = C.POINTER(C.c_int64)


class hppb200_stats(C.Structure):
    _fields_ = [("configs", C.c_uint64), ("pair_tests", C.c_uint64), ("gjk_iterations", C.c_uint64),
                ("edge_samples", C.c_uint64), ("kernel_launches", C.c_uint64), ("iter_cap_hits", C.c_uint64),
                ("last_kernel_ms", C.c_double), ("engine_ticks", C.c_uint64), ("discarded_tests", C.c_uint64)]


# every symbol include/hppb200.h declares (checked by tests/test_abi.py)
ABI_SYMBOLS = [
    "hppb200_abi_version", "hppb200_last_error", "hppb200_device_count", "hppb200_measure_fp64_peak", "hppb200_measure_tmem_read_peak",
    "hppb200_model_create", "hppb200_model_destroy", "hppb200_model_nq", "hppb200_model_nv",
    "hppb200_model_num_config_pairs", "hppb200_model_get_config_pairs", "hppb200_model_num_parameterized_pairs", "hppb200_model_num_body_pairs",
    "hppb200_model_get_body_pairs", "hppb200_model_num_body_geom_pairs", "hppb200_model_get_body_geom_pairs",
    "hppb200_model_num_coefficients", "hppb200_model_get_coefficients", "hppb200_model_get_body_radii",
    "hppb200_workspace_create", "hppb200_workspace_destroy", "hppb200_get_stats", "hppb200_workspace_set_stream",
    "hppb200_validate_configs", "hppb200_validate_configs_noalloc", "hppb200_geometry_placements", "hppb200_pair_distances", "hppb200_all_collisions", "hppb200_model_num_distance_pairs",
    "hppb200_distance_between_objects",
    "hppb200_straight_lengths", "hppb200_path_eval", "hppb200_validate_edges", "hppb200_validate_edges_indexed", "hppb200_validate_edge_chains", "hppb200_validate_spline_edges", "hppb200_workspace_set_max_iterations",
    "hppb200_multi_create", "hppb200_multi_destroy", "hppb200_multi_num_devices", "hppb200_multi_workspace",
    "hppb200_multi_validate_configs", "hppb200_multi_validate_edges",
    "hppb200_pack_mask", "hppb200_unpack_mask", "hppb200_knn_merge_shards",
    "hppb200_kprm_link_candidates", "hppb200_roadmap_merge_edges",
    "hppb200_exchange_create", "hppb200_exchange_open_peers", "hppb200_exchange_mask", "hppb200_exchange_status", "hppb200_exchange_destroy",
    "hppb200_knn_create", "hppb200_knn_destroy", "hppb200_knn_append", "hppb200_knn_relabel", "hppb200_knn_set_labels", "hppb200_knn_size",
    "hppb200_knn_query", "hppb200_knn_within_ball", "hppb200_knn_tc_stats", "hppb200_knn_tc_filter_time", "hppb200_knn_latency_time", "hppb200_knn_tc_debug_tile",
]

_LIB = None


def build(verbose=False):
    """Compile the CUDA extension in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    out = subprocess.run(["make", "-C", os.path.join(HERE, "csrc")], capture_output=True, text=True)
    if verbose or out.returncode != 0:
        print(out.stdout[-4000:], out.stderr[-4000:])
    if out.returncode != 0:
        raise RuntimeError("building libhppb200.so failed")


def lib():
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("CUDA extension %s is missing: run __graft_entry__.build() "
                               "(there is no CPU fallback)" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        L.hppb200_last_error.restype = C.c_char_p
        L.hppb200_knn_size.restype = C.c_size_t
        _LIB = L
    return _LIB


class EngineError(RuntimeError):
    pass


def _check(rc):
    if rc == OK:
        return
    msg = lib().hppb200_last_error().decode()
    if rc == ERR_INVALID_ARGUMENT:
        raise ValueError(msg)  # std::invalid_argument in the reference
    raise EngineError(msg)  # std::runtime_error in the reference


def _p(a, t):
    return None if a is None else a.ctypes.data_as(t)


def _is_torch(x):
    return type(x).__module__.startswith("torch")


def measure_fp64_peak(device=0):
    v = C.c_double()
    _check(lib().hppb200_measure_fp64_peak(C.c_int(device), C.byref(v)))
    return v.value


def measure_tmem_read_peak(device=0):
    v = C.c_double()
    _check(lib().hppb200_measure_tmem_read_peak(C.c_int(device), C.byref(v)))
    return v.value


def kprm_link_candidates(nbrs, dist, symmetric_distance):
    """hppb200_kprm_link_candidates: (kind [n][k] u8, mirror [n][k] i64) of the directed kPRM* candidates (host-side C++)."""
    nbrs = np.ascontiguousarray(nbrs, np.int64)
    n, k = nbrs.shape
    dist = np.ascontiguousarray(dist, np.float64) if dist is not None else None
    kind, mirror = np.zeros((n, k), np.uint8), np.zeros((n, k), np.int64)
    _check(lib().hppb200_kprm_link_candidates(C.c_size_t(n), C.c_int(k), nbrs.ctypes.data_as(C.c_void_p),
                                              dist.ctypes.data_as(C.c_void_p) if dist is not None else None,
                                              C.c_int(1 if symmetric_distance else 0), kind.ctypes.data_as(C.c_void_p),
                                              mirror.ctypes.data_as(C.c_void_p)))
    return kind, mirror


def roadmap_merge_edges(labels, pairs):
    """hppb200_roadmap_merge_edges: the labels after adding `pairs` ([m][2]) in order (the second component joins the first)."""
    labels = np.array(labels, np.int64, copy=True)
    pairs = np.ascontiguousarray(pairs, np.int64).reshape(-1, 2)
    _check(lib().hppb200_roadmap_merge_edges(C.c_size_t(len(labels)), labels.ctypes.data_as(C.c_void_p),
                                             pairs.ctypes.data_as(C.c_void_p), C.c_size_t(len(pairs))))
    return labels


class Model:
    """Compiled robot + obstacle tables (immutable, shareable between workspaces)."""

    def __init__(self, desc):
        self.desc = desc if isinstance(desc, ModelDescription) else ModelDescription.load(desc)
        self._struct = self.desc.struct()
        self.h = C.c_void_p()
        _check(lib().hppb200_model_create(C.byref(self._struct), C.byref(self.h)))
        self.nq = lib().hppb200_model_nq(self.h)
        self.nv = lib().hppb200_model_nv(self.h)

    def __del__(self):
        try:
            if getattr(self, "h", None):
                lib().hppb200_model_destroy(self.h)
                self.h = None
        except Exception:  # interpreter shutdown
            pass

    def config_pairs(self):
        n = lib().hppb200_model_num_config_pairs(self.h)
        a, b, m = np.zeros(n, np.int32), np.zeros(n, np.int32), np.zeros(n)
        lib().hppb200_model_get_config_pairs(self.h, _p(a, _ip), _p(b, _ip), _p(m, _dp))
        return a, b, m

    def body_pairs(self):
        n = lib().hppb200_model_num_body_pairs(self.h)
        ja, jb = np.zeros(n, np.int32), np.zeros(n, np.int32)
        pb, cb, rb = (np.zeros(n + 1, np.int32) for _ in range(3))
        lib().hppb200_model_get_body_pairs(self.h, _p(ja, _ip), _p(jb, _ip), _p(pb, _ip), _p(cb, _ip), _p(rb, _ip))
        ng = lib().hppb200_model_num_body_geom_pairs(self.h)
        ga, gb, gm = np.zeros(ng, np.int32), np.zeros(ng, np.int32), np.zeros(ng)
        lib().hppb200_model_get_body_geom_pairs(self.h, _p(ga, _ip), _p(gb, _ip), _p(gm, _dp))
        out = dict(joint_a=ja, joint_b=jb, pair_begin=pb, coef_begin=cb, rcoef_begin=rb, geomA=ga, geomB=gb, margin=gm)
        for rev, key in ((0, "coef"), (1, "rcoef")):
            nc = lib().hppb200_model_num_coefficients(self.h, rev)
            j, v = np.zeros(nc, np.int32), np.zeros(nc)
            lib().hppb200_model_get_coefficients(self.h, rev, _p(j, _ip), _p(v, _dp))
            out[key + "_joint"], out[key + "_value"] = j, v
        return out

    def body_radii(self):
        r = np.zeros(self.desc.njoints + 1)
        lib().hppb200_model_get_body_radii(self.h, _p(r, _dp))
        return r


class Workspace:
    """One CUDA stream + scratch on one device.  Accepts numpy arrays (host pointers, the
    library stages them) or CUDA torch tensors (device pointers, zero copy)."""

    def __init__(self, model, device=0):
        self.model = model
        self.device = device
        self.h = C.c_void_p()
        _check(lib().hppb200_workspace_create(model.h, C.c_int(device), C.byref(self.h)))

    def __del__(self):
        try:
            if getattr(self, "h", None):
                lib().hppb200_workspace_destroy(self.h)
                self.h = None
        except Exception:  # interpreter shutdown
            pass

    def set_stream(self, cuda_stream_ptr):
        _check(lib().hppb200_workspace_set_stream(self.h, C.c_void_p(cuda_stream_ptr)))

    def set_max_iterations(self, n):
        _check(lib().hppb200_workspace_set_max_iterations(self.h, C.c_uint32(n)))

    def stats(self):
        s = hppb200_stats()
        _check(lib().hppb200_get_stats(self.h, C.byref(s)))
        return {k: getattr(s, k) for k, _ in s._fields_}

    # -- raw pointer forms (host or device pointers, as ints) -----------------------------
    def validate_configs_ptr(self, q_ptr, n, valid_ptr, first_ptr=0, reason_ptr=0, flags=0):
        _check(lib().hppb200_validate_configs(self.h, C.c_void_p(q_ptr), C.c_size_t(n), C.c_uint32(flags),
                                              C.c_void_p(valid_ptr), C.c_void_p(first_ptr or None),
                                              C.c_void_p(reason_ptr or None)))

    def validate_configs_noalloc_ptr(self, q_ptr, n, valid_ptr, first_ptr, reason_ptr, flags=DEVICE_POINTERS):
        """hppb200_validate_configs_noalloc: device pointers only, scratch in the library's static device array."""
        _check(lib().hppb200_validate_configs_noalloc(self.h, C.c_void_p(q_ptr), C.c_size_t(n), C.c_uint32(flags),
                                                      C.c_void_p(valid_ptr or None), C.c_void_p(first_ptr or None),
                                                      C.c_void_p(reason_ptr or None)))

    def validate_edges_ptr(self, method, param, reverse, q1_ptr, q2_ptr, L_ptr, n, valid_ptr, tfail_ptr=0,
                           tlast_ptr=0, pair_ptr=0, reason_ptr=0, flags=0):
        _check(lib().hppb200_validate_edges(self.h, C.c_int(method), C.c_double(param), C.c_int(1 if reverse else 0),
                                            C.c_void_p(q1_ptr), C.c_void_p(q2_ptr), C.c_void_p(L_ptr), C.c_size_t(n),
                                            C.c_uint32(flags), C.c_void_p(valid_ptr), C.c_void_p(tfail_ptr or None),
                                            C.c_void_p(tlast_ptr or None), C.c_void_p(pair_ptr or None),
                                            C.c_void_p(reason_ptr or None)))

    # -- numpy forms ------------------------------------------------------------------------
    def validate_configs(self, q, check_bounds=False):
        q = np.ascontiguousarray(q, np.float64).reshape(-1, self.model.nq)
        n = q.shape[0]
        valid, first, reason = np.zeros(n, np.uint8), np.full(n, -1, np.int32), np.zeros(n, np.uint8)
        self.validate_configs_ptr(q.ctypes.data, n, valid.ctypes.data, first.ctypes.data, reason.ctypes.data,
                                  CHECK_JOINT_BOUNDS if check_bounds else 0)
        return valid, first, reason

    def geometry_placements(self, q):
        q = np.ascontiguousarray(q, np.float64).reshape(-1, self.model.nq)
        out = np.zeros((q.shape[0], self.model.desc.ngeoms, 12))
        _check(lib().hppb200_geometry_placements(self.h, _p(q, _dp), C.c_size_t(q.shape[0]), C.c_uint32(0), _p(out, _dp)))
        return out

    def pair_distances(self, q):
        q = np.ascontiguousarray(q, np.float64).reshape(-1, self.model.nq)
        npairs = lib().hppb200_model_num_config_pairs(self.model.h)
        out = np.zeros((q.shape[0], npairs))
        _check(lib().hppb200_pair_distances(self.h, _p(q, _dp), C.c_size_t(q.shape[0]), C.c_uint32(0), _p(out, _dp)))
        return out

    def all_collisions(self, q):
        """CollisionValidation with computeAllContacts: (valid, first, list of colliding pair indices per configuration)"""
        q = np.ascontiguousarray(q, np.float64).reshape(-1, self.model.nq)
        n = q.shape[0]
        npairs = lib().hppb200_model_num_config_pairs(self.model.h)
        words = (npairs + 31) // 32
        valid, first, count = np.zeros(n, np.uint8), np.full(n, -1, np.int32), np.zeros(n, np.int32)
        mask = np.zeros((n, words), np.uint32)
        _check(lib().hppb200_all_collisions(self.h, _p(q, _dp), C.c_size_t(n), C.c_uint32(0), _p(valid, _u8p), _p(first, _ip),
                                            mask.ctypes.data_as(C.c_void_p), _p(count, _ip)))
        bits = np.unpackbits(mask.view(np.uint8), axis=1, bitorder="little")[:, :npairs]
        return valid, first, [np.nonzero(b)[0] for b in bits], count

    def distance_between_objects(self, q):
        q = np.ascontiguousarray(q, np.float64).reshape(-1, self.model.nq)
        per = lib().hppb200_model_num_distance_pairs(self.model.h)
        out = np.zeros((q.shape[0], per))
        _check(lib().hppb200_distance_between_objects(self.h, _p(q, _dp), C.c_size_t(q.shape[0]), C.c_uint32(0), _p(out, _dp)))
        return out

    def straight_lengths(self, q1, q2):
        q1 = np.ascontiguousarray(q1, np.float64).reshape(-1, self.model.nq)
        q2 = np.ascontiguousarray(q2, np.float64).reshape(-1, self.model.nq)
        L = np.zeros(q1.shape[0])
        _check(lib().hppb200_straight_lengths(self.h, _p(q1, _dp), _p(q2, _dp), C.c_size_t(q1.shape[0]), C.c_uint32(0),
                                              _p(L, _dp)))
        return L

    def path_eval(self, q1, q2, L, t):
        """StraightPath::impl_compute for many edges: the configuration of edge i at parameter t[i]."""
        q1 = np.ascontiguousarray(q1, np.float64).reshape(-1, self.model.nq)
        q2 = np.ascontiguousarray(q2, np.float64).reshape(-1, self.model.nq)
        L, t = np.ascontiguousarray(L, np.float64), np.ascontiguousarray(t, np.float64)
        q = np.zeros_like(q1)
        _check(lib().hppb200_path_eval(self.h, _p(q1, _dp), _p(q2, _dp), _p(L, _dp), _p(t, _dp), C.c_size_t(q1.shape[0]),
                                       C.c_uint32(0), _p(q, _dp)))
        return q

    def validate_edges(self, method, param, q1, q2, L, reverse=False, check_bounds=False):
        q1 = np.ascontiguousarray(q1, np.float64).reshape(-1, self.model.nq)
        q2 = np.ascontiguousarray(q2, np.float64).reshape(-1, self.model.nq)
        L = np.ascontiguousarray(L, np.float64)
        n = q1.shape[0]
        valid, reason = np.zeros(n, np.uint8), np.zeros(n, np.uint8)
        t_fail, t_last = np.zeros(n), np.zeros(n)
        pair = np.full(n, -1, np.int32)
        self.validate_edges_ptr(method, param, reverse, q1.ctypes.data, q2.ctypes.data, L.ctypes.data, n,
                                valid.ctypes.data, t_fail.ctypes.data, t_last.ctypes.data, pair.ctypes.data,
                                reason.ctypes.data, CHECK_JOINT_BOUNDS if check_bounds else 0)
        return dict(valid=valid, t_fail=t_fail, t_last_valid=t_last, pair=pair, reason=reason)


    def validate_edges_indexed(self, nn, method, param, node_a, node_b, reverse=False):
        """edges between nodes of the device roadmap mirror `nn`, given by node index"""
        a = np.ascontiguousarray(node_a, np.int64)
        b = np.ascontiguousarray(node_b, np.int64)
        n = a.shape[0]
        valid, reason = np.zeros(n, np.uint8), np.zeros(n, np.uint8)
        t_fail, t_last, L = np.zeros(n), np.zeros(n), np.zeros(n)
        pair = np.full(n, -1, np.int32)
        _check(lib().hppb200_validate_edges_indexed(self.h, nn.h, C.c_int(method), C.c_double(param), C.c_int(1 if reverse else 0),
                                                    _p(a, _i64p), _p(b, _i64p), C.c_size_t(n), C.c_uint32(0), _p(valid, _u8p),
                                                    _p(t_fail, _dp), _p(t_last, _dp), _p(pair, _ip), _p(reason, _u8p), _p(L, _dp)))
        return dict(valid=valid, t_fail=t_fail, t_last_valid=t_last, pair=pair, reason=reason, length=L)

    def validate_edge_chains(self, method, param, q1, q2, L, chain_begin, reverse=False):
        """PathVectors of StraightPaths (ContinuousValidation::validate recursion, src/continuous-validation.cc:153-192)."""
        q1 = np.ascontiguousarray(q1, np.float64).reshape(-1, self.model.nq)
        q2 = np.ascontiguousarray(q2, np.float64).reshape(-1, self.model.nq)
        L = np.ascontiguousarray(L, np.float64)
        cb = np.ascontiguousarray(chain_begin, np.int32)
        nc = cb.shape[0] - 1
        valid, reason = np.zeros(nc, np.uint8), np.zeros(nc, np.uint8)
        t_fail, t_last = np.zeros(nc), np.zeros(nc)
        pair, bad = np.full(nc, -1, np.int32), np.full(nc, -1, np.int32)
        _check(lib().hppb200_validate_edge_chains(self.h, C.c_int(method), C.c_double(param), C.c_int(1 if reverse else 0),
                                                  _p(q1, _dp), _p(q2, _dp), _p(L, _dp), C.c_size_t(q1.shape[0]), _p(cb, _ip),
                                                  C.c_size_t(nc), C.c_uint32(0), _p(valid, _u8p), _p(t_fail, _dp), _p(t_last, _dp),
                                                  _p(pair, _ip), _p(reason, _u8p), _p(bad, _ip)))
        return dict(valid=valid, t_fail=t_fail, t_last_valid=t_last, pair=pair, reason=reason, first_bad_edge=bad)

    def validate_spline_edges(self, method, param, base, params, L, reverse=False, check_bounds=False):
        """edges that are path::Spline<BernsteinBasis, 3> (tests/test-continuous-validation.cc:288-427):
        base [n][nq], params [n][4][nv], L [n]; same outputs as validate_edges"""
        nq = self.model.nq
        base = np.ascontiguousarray(base, np.float64).reshape(-1, nq)
        params = np.ascontiguousarray(params, np.float64).reshape(base.shape[0], 4, nq)
        L = np.ascontiguousarray(L, np.float64)
        n = base.shape[0]
        valid, reason = np.zeros(n, np.uint8), np.zeros(n, np.uint8)
        t_fail, t_last = np.zeros(n), np.zeros(n)
        pair = np.full(n, -1, np.int32)
        _check(lib().hppb200_validate_spline_edges(self.h, C.c_int(method), C.c_double(param), C.c_int(1 if reverse else 0),
                                                   _p(base, _dp), _p(params, _dp), _p(L, _dp), C.c_size_t(n),
                                                   C.c_uint32(CHECK_JOINT_BOUNDS if check_bounds else 0), _p(valid, _u8p),
                                                   _p(t_fail, _dp), _p(t_last, _dp), _p(pair, _ip), _p(reason, _u8p)))
        return dict(valid=valid, t_fail=t_fail, t_last_valid=t_last, pair=pair, reason=reason)


def spline_steer(q1, v1, q2, v2, L):
    """steeringMethod::Spline<BernsteinBasis, 3>::steer(q1, {1}, v1, q2, {1}, v2, L) (reference
    src/steering-method/spline.cc:75-142) for vector-space configurations, in closed form: the cubic Bezier curve over
    [0, L] from q1 to q2 with end velocities v1, v2 (base = q1; control points 0, L v1 / 3, (q2 - q1) - L v2 / 3, q2 - q1).
    The reference solves the same 4 x 4 system with a Jacobi SVD: equal up to rounding.  Returns (base, params [.., 4, nv])."""
    q1, v1, q2, v2 = (np.asarray(x, np.float64) for x in (q1, v1, q2, v2))
    L = np.asarray(L, np.float64)[..., None]
    d = q2 - q1
    return q1.copy(), np.stack([np.zeros_like(d), L * v1 / 3.0, d - L * v2 / 3.0, d], axis=-2)


class MultiGpu:
    """hppb200_multi_*: a device list, sharded internally (host arrays in, host arrays out)."""

    def __init__(self, model, devices=None):
        self.model = model
        self.h = C.c_void_p()
        dv = None if not devices else (C.c_int * len(devices))(*devices)
        _check(lib().hppb200_multi_create(model.h, dv, C.c_int(len(devices) if devices else 0), C.byref(self.h)))
        self.ndevices = lib().hppb200_multi_num_devices(self.h)

    def __del__(self):
        try:
            if getattr(self, "h", None):
                lib().hppb200_multi_destroy(self.h)
                self.h = None
        except Exception:  # interpreter shutdown
            pass

    def shard_stats(self, rank):
        lib().hppb200_multi_workspace.restype = C.c_void_p
        w = lib().hppb200_multi_workspace(self.h, C.c_int(rank))
        s = hppb200_stats()
        _check(lib().hppb200_get_stats(C.c_void_p(w), C.byref(s)))
        return {k: getattr(s, k) for k, _ in s._fields_}

    def validate_configs(self, q, check_bounds=False):
        q = np.ascontiguousarray(q, np.float64).reshape(-1, self.model.nq)
        n = q.shape[0]
        valid, first, reason = np.zeros(n, np.uint8), np.full(n, -1, np.int32), np.zeros(n, np.uint8)
        _check(lib().hppb200_multi_validate_configs(self.h, _p(q, _dp), C.c_size_t(n), C.c_uint32(CHECK_JOINT_BOUNDS if check_bounds else 0),
                                                    _p(valid, _u8p), _p(first, _ip), _p(reason, _u8p)))
        return valid, first, reason

    def validate_edges(self, method, param, q1, q2, L, reverse=False):
        q1 = np.ascontiguousarray(q1, np.float64).reshape(-1, self.model.nq)
        q2 = np.ascontiguousarray(q2, np.float64).reshape(-1, self.model.nq)
        L = np.ascontiguousarray(L, np.float64)
        n = q1.shape[0]
        valid, reason = np.zeros(n, np.uint8), np.zeros(n, np.uint8)
        t_fail, t_last = np.zeros(n), np.zeros(n)
        pair = np.full(n, -1, np.int32)
        _check(lib().hppb200_multi_validate_edges(self.h, C.c_int(method), C.c_double(param), C.c_int(1 if reverse else 0), _p(q1, _dp),
                                                  _p(q2, _dp), _p(L, _dp), C.c_size_t(n), C.c_uint32(0), _p(valid, _u8p), _p(t_fail, _dp),
                                                  _p(t_last, _dp), _p(pair, _ip), _p(reason, _u8p)))
        return dict(valid=valid, t_fail=t_fail, t_last_valid=t_last, pair=pair, reason=reason)


class NearestNeighbor:
    """Device mirror of the roadmap nodes + brute-force WeighedDistance search
    (nearestNeighbor::Basic, reference src/nearest-neighbor/basic.cc)."""

    def __init__(self, model, device=0):
        self.model = model
        self.h = C.c_void_p()
        _check(lib().hppb200_knn_create(model.h, C.c_int(device), C.byref(self.h)))

    def __del__(self):
        try:
            if getattr(self, "h", None):
                lib().hppb200_knn_destroy(self.h)
                self.h = None
        except Exception:  # interpreter shutdown
            pass

    def __len__(self):
        return int(lib().hppb200_knn_size(self.h))

    def append(self, q, cc=None):
        q = np.ascontiguousarray(q, np.float64).reshape(-1, self.model.nq)
        cc = None if cc is None else np.ascontiguousarray(cc, np.int32)
        _check(lib().hppb200_knn_append(self.h, _p(q, _dp), _p(cc, _ip), C.c_size_t(q.shape[0])))

    def relabel(self, from_cc, to_cc):
        _check(lib().hppb200_knn_relabel(self.h, C.c_int32(from_cc), C.c_int32(to_cc)))

    def set_labels(self, cc):
        cc = np.ascontiguousarray(cc, np.int32)
        _check(lib().hppb200_knn_set_labels(self.h, _p(cc, _ip), C.c_size_t(cc.shape[0])))

    def query(self, queries, k=1, cc=-1):
        queries = np.ascontiguousarray(queries, np.float64).reshape(-1, self.model.nq)
        n = queries.shape[0]
        idx, dist = np.zeros((n, k), np.int64), np.zeros((n, k))
        _check(lib().hppb200_knn_query(self.h, _p(queries, _dp), C.c_size_t(n), C.c_int(k), C.c_int32(cc),
                                       C.c_uint32(0), _p(idx, _i64p), _p(dist, _dp)))
        return idx, dist

    def query_ptr(self, q_ptr, n, k, idx_ptr, dist_ptr, cc=-1):
        """device pointers: queries [n][nq] f64, idx [n][k] i64, dist [n][k] f64 (synchronises the mirror's stream)"""
        _check(lib().hppb200_knn_query(self.h, C.c_void_p(q_ptr), C.c_size_t(n), C.c_int(k), C.c_int32(cc), C.c_uint32(DEVICE_POINTERS),
                                       C.c_void_p(idx_ptr), C.c_void_p(dist_ptr)))

    def tc_stats(self):
        """(used, stages, queries redone exactly, candidates of the last stage) of the last query."""
        out = np.zeros(4, np.int64)
        _check(lib().hppb200_knn_tc_stats(self.h, _p(out, _i64p)))
        ms, pairs = C.c_double(0), C.c_double(0)
        _check(lib().hppb200_knn_tc_filter_time(self.h, C.byref(ms), C.byref(pairs)))
        return dict(used=bool(out[0]), stages=int(out[1]), redone=int(out[2]), candidates=int(out[3]),
                    filter_ms=ms.value, filter_pairs=pairs.value)

    def latency_time(self):
        """(ms, bytes scanned) of the single-query kernel of the last query (mirror created under HPPB200_KNN_TIMING=1)"""
        ms, nbytes = C.c_double(0), C.c_double(0)
        _check(lib().hppb200_knn_latency_time(self.h, C.byref(ms), C.byref(nbytes)))
        return ms.value, nbytes.value

    def tc_debug_tile(self, queries, swap_offsets=0):
        queries = np.ascontiguousarray(queries, np.float64).reshape(-1, self.model.nq)
        out = np.zeros((128, 256), np.float32)
        _check(lib().hppb200_knn_tc_debug_tile(self.h, _p(queries, _dp), C.c_size_t(queries.shape[0]),
                                                 C.c_int(swap_offsets), out.ctypes.data_as(C.c_void_p)))
        return out

    def within_ball(self, queries, radius, cc=-1, cap=4096):
        queries = np.ascontiguousarray(queries, np.float64).reshape(-1, self.model.nq)
        n = queries.shape[0]
        idx, count = np.full((n, cap), -1, np.int64), np.zeros(n, np.int64)
        _check(lib().hppb200_knn_within_ball(self.h, _p(queries, _dp), C.c_size_t(n), C.c_double(radius),
                                             C.c_int32(cc), C.c_size_t(cap), _p(idx, _i64p), _p(count, _i64p)))
        return [idx[i, :min(count[i], cap)].copy() for i in range(n)], count


from .planners import (BiRRTPlanner, DiffusingPlanner, EngineBackend, KPrmStar, PartialShortcut, RandomShortcut,  # noqa: E402,F401
                       Roadmap, SimpleShortcut, VisibilityPrmPlanner)  # noqa: E402,F401
from .validators import (CollisionValidation, ConfigValidations, Dichotomy, Discretized,  # noqa: E402,F401
                         JointBoundValidation, Progressive, StraightPath, ValidationReport)
