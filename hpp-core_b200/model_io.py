"""Model description files <-> `hppb200_model_desc` (include/hppb200.h).

Pure data plumbing: loads models/*.json into numpy arrays and exposes them as the ctypes
struct the C ABI takes.  No validation arithmetic happens here.
"""
import ctypes as C
import json
import os

import numpy as np

c_double_p = C.POINTER(C.c_double)
c_int32_p = C.POINTER(C.c_int32)


class hppb200_pair_op(C.Structure):
    _fields_ = [("kind", C.c_int32), ("a", C.c_int32), ("b", C.c_int32), ("flag", C.c_int32), ("value", C.c_double),
                ("data", C.c_void_p)]


OP_ADD_OBSTACLE_TO_JOINT, OP_REMOVE_OBSTACLE_FROM_JOINT, OP_FILTER_COLLISION_PAIRS, OP_SET_SECURITY_MARGINS, \
    OP_SET_MARGIN_BETWEEN_BODIES = 1, 2, 3, 4, 5
OP_ADD_OBSTACLE = 6


class hppb200_model_desc(C.Structure):
    _fields_ = [
        ("njoints", C.c_int32),
        ("joint_parent", c_int32_p),
        ("joint_type", c_int32_p),
        ("joint_placement", c_double_p),
        ("q_lower", c_double_p),
        ("q_upper", c_double_p),
        ("ngeoms", C.c_int32),
        ("geom_joint", c_int32_p),
        ("geom_placement", c_double_p),
        ("geom_shape", c_int32_p),
        ("geom_params", c_double_p),
        ("geom_vert_begin", c_int32_p),
        ("geom_vert_count", c_int32_p),
        ("nself_pairs", C.c_int32),
        ("self_pairs", c_int32_p),
        ("nobstacles", C.c_int32),
        ("obst_placement", c_double_p),
        ("obst_shape", c_int32_p),
        ("obst_params", c_double_p),
        ("obst_vert_begin", c_int32_p),
        ("obst_vert_count", c_int32_p),
        ("nvertices", C.c_int32),
        ("vertices", c_double_p),
        ("security_margins", c_double_p),
        ("break_distance", C.c_double),
        ("weights", c_double_p),
        ("obst_global", C.POINTER(C.c_uint8)),
        ("npair_ops", C.c_int32),
        ("pair_ops", C.POINTER(hppb200_pair_op)),
        ("check_parameterized", C.c_int32),
    ]


def _f64(x, shape=None):
    a = np.ascontiguousarray(np.asarray(x, dtype=np.float64))
    return a.reshape(shape) if shape is not None else a


def _i32(x, shape=None):
    a = np.ascontiguousarray(np.asarray(x, dtype=np.int32))
    return a.reshape(shape) if shape is not None else a


class ModelDescription:
    """Owns the arrays behind a `hppb200_model_desc` (keep it alive while the struct is used)."""

    def __init__(self, d):
        self.name = d.get("name", "model")
        nj = len(d["joint_parent"])
        self.njoints = nj
        self.joint_parent = _i32(d["joint_parent"])
        self.joint_type = _i32(d["joint_type"])
        self.joint_placement = _f64(d["joint_placement"], (nj, 12))
        self.q_lower = _f64(d["q_lower"])
        self.q_upper = _f64(d["q_upper"])
        self.nq = int(self.q_lower.shape[0])
        ng = len(d["geom_joint"])
        self.ngeoms = ng
        self.geom_joint = _i32(d["geom_joint"])
        self.geom_placement = _f64(d["geom_placement"], (ng, 12))
        self.geom_shape = _i32(d["geom_shape"])
        self.geom_params = _f64(d["geom_params"], (ng, 4))
        self.geom_vert_begin = _i32(d.get("geom_vert_begin", [0] * ng))
        self.geom_vert_count = _i32(d.get("geom_vert_count", [0] * ng))
        self.self_pairs = _i32(d.get("self_pairs", []), (-1, 2))
        no = len(d.get("obst_shape", []))
        self.nobstacles = no
        self.obst_placement = _f64(d.get("obst_placement", []), (no, 12))
        self.obst_shape = _i32(d.get("obst_shape", []))
        self.obst_params = _f64(d.get("obst_params", []), (no, 4))
        self.obst_vert_begin = _i32(d.get("obst_vert_begin", [0] * no))
        self.obst_vert_count = _i32(d.get("obst_vert_count", [0] * no))
        self.vertices = _f64(d.get("vertices", []), (-1, 3))
        sm = d.get("security_margins")
        self.security_margins = None if sm is None else _f64(sm, (nj + 1, nj + 1))
        self.break_distance = float(d.get("break_distance", 1e-2))
        w = d.get("weights")
        self.weights = None if w is None else _f64(w)
        self.geom_names = d.get("geom_names")
        self.obst_names = d.get("obst_names")
        # pair-table edits (ObstacleUserInterface call log): list of (kind, a, b, flag, value, data) with data a matrix
        # ((njoints+1)^2 int32 for OP_FILTER_COLLISION_PAIRS, float64 for OP_SET_SECURITY_MARGINS) or None
        og = d.get("obst_global")
        self.obst_global = None if og is None else np.ascontiguousarray(np.asarray(og, dtype=np.uint8))
        self.pair_ops = [tuple(op) for op in d.get("pair_ops", [])]
        self.check_parameterized = bool(d.get("check_parameterized", False))
        self._op_keep = []

    @classmethod
    def load(cls, path):
        if not os.path.isabs(path) and not os.path.exists(path):
            path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "models", path)
        with open(path) as f:
            return cls(json.load(f))

    def struct(self):
        def dp(a):
            return None if a is None or a.size == 0 else a.ctypes.data_as(c_double_p)

        def ip(a):
            return None if a is None or a.size == 0 else a.ctypes.data_as(c_int32_p)

        s = hppb200_model_desc()
        s.njoints = self.njoints
        s.joint_parent, s.joint_type = ip(self.joint_parent), ip(self.joint_type)
        s.joint_placement = dp(self.joint_placement)
        s.q_lower, s.q_upper = dp(self.q_lower), dp(self.q_upper)
        s.ngeoms = self.ngeoms
        s.geom_joint, s.geom_placement = ip(self.geom_joint), dp(self.geom_placement)
        s.geom_shape, s.geom_params = ip(self.geom_shape), dp(self.geom_params)
        s.geom_vert_begin, s.geom_vert_count = ip(self.geom_vert_begin), ip(self.geom_vert_count)
        s.nself_pairs = int(self.self_pairs.shape[0])
        s.self_pairs = ip(self.self_pairs)
        s.nobstacles = self.nobstacles
        s.obst_placement, s.obst_shape = dp(self.obst_placement), ip(self.obst_shape)
        s.obst_params = dp(self.obst_params)
        s.obst_vert_begin, s.obst_vert_count = ip(self.obst_vert_begin), ip(self.obst_vert_count)
        s.nvertices = int(self.vertices.shape[0])
        s.vertices = dp(self.vertices)
        s.security_margins = dp(self.security_margins)
        s.break_distance = self.break_distance
        s.weights = dp(self.weights)
        if self.obst_global is not None and self.obst_global.size:
            s.obst_global = self.obst_global.ctypes.data_as(C.POINTER(C.c_uint8))
        s.check_parameterized = 1 if self.check_parameterized else 0
        if self.pair_ops:
            n1 = self.njoints + 1
            arr = (hppb200_pair_op * len(self.pair_ops))()
            keep = []
            for i, op in enumerate(self.pair_ops):
                kind, a, b, flag, value, data = (list(op) + [0, 0, 0, 0.0, None])[:6]
                arr[i].kind, arr[i].a, arr[i].b, arr[i].flag, arr[i].value = int(kind), int(a), int(b), int(flag), float(value)
                if data is not None:
                    m = np.ascontiguousarray(np.asarray(data, dtype=np.int32 if kind == OP_FILTER_COLLISION_PAIRS else np.float64)
                                             ).reshape(n1, n1)
                    keep.append(m)
                    arr[i].data = m.ctypes.data
            self._op_keep = [arr, keep]  # owned as long as the description
            s.npair_ops = len(self.pair_ops)
            s.pair_ops = C.cast(arr, C.POINTER(hppb200_pair_op))
        return s


def splitmix64(x):
    """Counter-based RNG shared by host inputs and the device generator."""
    x = (np.asarray(x, dtype=np.uint64) + np.uint64(0x9E3779B97F4A7C15))
    z = x.copy()
    z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return z ^ (z >> np.uint64(31))


def uniform_configs(model, n, seed=12345, offset=0):
    """n configurations uniform in the joint limits: u = (splitmix64(seed, index) >> 11) * 2^-53
    (configurationShooter::Uniform semantics, reference src/configuration-shooter/uniform.cc:38-61)."""
    nq = model.nq
    idx = (np.arange(offset * nq, (offset + n) * nq, dtype=np.uint64)
           + np.uint64(seed) * np.uint64(0x100000001B3))
    with np.errstate(over="ignore"):
        bits = splitmix64(idx)
    u = (bits >> np.uint64(11)).astype(np.float64) * (2.0 ** -53)
    u = u.reshape(n, nq)
    return normalize_quaternions(model, model.q_lower[None, :] + u * (model.q_upper - model.q_lower)[None, :])


def freeflyer_ranks(model):
    """Configuration rank of every free-flyer joint (7 entries each: x y z, qx qy qz qw)."""
    ranks, r = [], 0
    for t in model.joint_type:
        if int(t) == 6:
            ranks.append(r)
        r += 7 if int(t) == 6 else 1
    return ranks


def normalize_quaternions(model, q):
    """Scales the quaternion entries of free-flyer joints to unit norm (input data preparation)."""
    q = np.array(q, dtype=np.float64, copy=True)
    for r in freeflyer_ranks(model):
        q[..., r + 3:r + 7] /= np.linalg.norm(q[..., r + 3:r + 7], axis=-1, keepdims=True)
    return q
