#!/usr/bin/env python3
"""Generate the robot/scene description files used by tests and bench.

Neither example-robot-data (UR5 URDF/SRDF/meshes) nor a URDF parser exists in this image, so the
UR5 is described here directly: kinematics from the ROS-Industrial ur_description URDF
(SURVEY.md Appendix A.1), collision geometry as capsule/box primitives (the real URDF uses STL
meshes -> coal BVH, not reproducible without coal; stated in DESIGN.md).  The JSON layout mirrors
`hppb200_model_desc` (include/hppb200.h).

    python models/make_models.py        # rewrites models/*.json deterministically
"""
import json
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))

REV_X, REV_Y, REV_Z, PRI_X, PRI_Y, PRI_Z, FREEFLYER = range(7)
SPHERE, CAPSULE, BOX, CONVEX = range(4)


def rpy(r, p, y):
    cr, sr, cp, sp, cy, sy = np.cos(r), np.sin(r), np.cos(p), np.sin(p), np.cos(y), np.sin(y)
    Rx = np.array([[1, 0, 0], [0, cr, -sr], [0, sr, cr]])
    Ry = np.array([[cp, 0, sp], [0, 1, 0], [-sp, 0, cp]])
    Rz = np.array([[cy, -sy, 0], [sy, cy, 0], [0, 0, 1]])
    R = Rz @ Ry @ Rx
    R[np.abs(R) < 1e-15] = 0.0  # exact zeros for the pi/2 rotations
    return R


def se3(R=None, t=(0, 0, 0)):
    R = np.eye(3) if R is None else np.asarray(R, float)
    return [float(x) for x in R.reshape(-1)] + [float(x) for x in t]


ALIGN_Z_TO_Y = rpy(-np.pi / 2, 0, 0)  # maps local z onto +y


def ur5():
    joints = [  # name, parent, type, xyz, rpy
        ("shoulder_pan_joint", 0, REV_Z, (0, 0, 0.089159), (0, 0, 0)),
        ("shoulder_lift_joint", 1, REV_Y, (0, 0.13585, 0), (0, np.pi / 2, 0)),
        ("elbow_joint", 2, REV_Y, (0, -0.1197, 0.425), (0, 0, 0)),
        ("wrist_1_joint", 3, REV_Y, (0, 0, 0.39225), (0, np.pi / 2, 0)),
        ("wrist_2_joint", 4, REV_Z, (0, 0.093, 0), (0, 0, 0)),
        ("wrist_3_joint", 5, REV_Y, (0, 0, 0.09465), (0, 0, 0)),
    ]
    geoms = [  # name, joint, shape, params, placement
        ("base_link_0", 0, CAPSULE, (0.075, 0.02), se3(t=(0, 0, 0.02))),
        ("shoulder_link_0", 1, CAPSULE, (0.07, 0.04), se3(ALIGN_Z_TO_Y, (0, 0.02, 0))),
        ("upper_arm_link_0", 2, CAPSULE, (0.055, 0.2125), se3(t=(0, 0, 0.2125))),
        ("forearm_link_0", 3, CAPSULE, (0.045, 0.196), se3(t=(0, 0, 0.196))),
        ("wrist_1_link_0", 4, CAPSULE, (0.04, 0.035), se3(ALIGN_Z_TO_Y, (0, 0.06, 0))),
        ("wrist_2_link_0", 5, CAPSULE, (0.04, 0.03), se3(t=(0, 0, 0.06))),
        ("wrist_3_link_0", 6, CAPSULE, (0.035, 0.02), se3(ALIGN_Z_TO_Y, (0, 0.04, 0))),
        ("tool_0", 6, BOX, (0.03, 0.02, 0.03), se3(t=(0, 0.0823 + 0.02, 0))),
    ]
    # self-collision pairs: every pair of geometries on non-adjacent links (what the SRDF leaves on)
    link_of = [g[1] for g in geoms]
    pairs = []
    for a in range(len(geoms)):
        for b in range(a + 1, len(geoms)):
            if abs(link_of[a] - link_of[b]) >= 2:
                pairs.append([a, b])
    return {
        "name": "ur5_primitives",
        "joint_names": [j[0] for j in joints],
        "joint_parent": [j[1] for j in joints],
        "joint_type": [j[2] for j in joints],
        "joint_placement": [se3(rpy(*j[4]), j[3]) for j in joints],
        "q_lower": [-np.pi] * 6,
        "q_upper": [np.pi] * 6,
        "geom_names": [g[0] for g in geoms],
        "geom_joint": [g[1] for g in geoms],
        "geom_shape": [g[2] for g in geoms],
        "geom_params": [list(g[3]) + [0.0] * (4 - len(g[3])) for g in geoms],
        "geom_placement": [g[4] for g in geoms],
        "geom_vert_begin": [0] * len(geoms),
        "geom_vert_count": [0] * len(geoms),
        "self_pairs": pairs,
        "vertices": [],
        "break_distance": 1e-2,
    }


def random_rotation(rng):
    q = rng.normal(size=4)
    q /= np.linalg.norm(q)
    x, y, z, w = q
    return np.array([
        [1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
        [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
        [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)],
    ])


def boxes(n, seed=20240607):
    """n box obstacles: sides U[0.1,0.4] m, centres in the shell 0.3 <= |xy| <= 0.9, z in [0,0.8]."""
    rng = np.random.default_rng(seed)
    out = {"obst_names": [], "obst_placement": [], "obst_shape": [], "obst_params": [],
           "obst_vert_begin": [], "obst_vert_count": []}
    for k in range(n):
        side = rng.uniform(0.1, 0.4, size=3)
        rad = rng.uniform(0.3, 0.9)
        ang = rng.uniform(-np.pi, np.pi)
        c = (rad * np.cos(ang), rad * np.sin(ang), rng.uniform(0.0, 0.8))
        out["obst_names"].append("box_%d" % k)
        out["obst_placement"].append(se3(random_rotation(rng), c))
        out["obst_shape"].append(BOX)
        out["obst_params"].append([float(s) / 2 for s in side] + [0.0])
        out["obst_vert_begin"].append(0)
        out["obst_vert_count"].append(0)
    return out


def fk_neutral(model):
    """oMi at q = 0 for vector-space models (plain numpy; data preparation only)."""
    n = len(model["joint_parent"])
    oMi = [(np.eye(3), np.zeros(3))]
    for i in range(n):
        P = np.array(model["joint_placement"][i])
        R, t = P[:9].reshape(3, 3), P[9:]
        Rp, tp = oMi[model["joint_parent"][i]]
        oMi.append((Rp @ R, Rp @ t + tp))
    return oMi


def body_radii(model):
    n = len(model["joint_parent"])
    r = np.zeros(n + 1)
    for g, j in enumerate(model["geom_joint"]):
        p = model["geom_params"][g]
        s = model["geom_shape"][g]
        if s == SPHERE:
            h = np.array([p[0]] * 3)
        elif s == CAPSULE:
            h = np.array([p[0], p[0], p[0] + p[1]])
        else:
            h = np.array(p[:3])
        P = np.array(model["geom_placement"][g])
        R, t = P[:9].reshape(3, 3), P[9:]
        for c in range(8):
            corner = np.array([h[0] if c & 1 else -h[0], h[1] if c & 2 else -h[1], h[2] if c & 4 else -h[2]])
            r[j] = max(r[j], np.linalg.norm(R @ corner + t))
    return r


def weighed_distance_weights(model):
    """WeighedDistance::computeWeights restated for revolute chains/trees at the neutral
    configuration (reference src/weighed-distance.cc:94-122,218-251), including the
    `radius + squaredNorm(dt)` quirk (:116-118) and the zero-weight replacement (:242-248)."""
    n = len(model["joint_parent"])
    oMi = fk_neutral(model)
    radii = body_radii(model)
    parent = [0] + list(model["joint_parent"])
    w = np.zeros(n)
    min_len = np.inf
    for i in range(1, n + 1):
        jt = model["joint_type"][i - 1]
        axis = oMi[i][0][:, jt % 3]
        length = 0.0
        for j in range(i, n + 1):
            k = j  # is j in the subtree of i ?
            while k > i:
                k = parent[k]
            if k != i:
                continue
            dt = oMi[j][1] - oMi[i][1]
            if jt <= REV_Z:
                lin, ang = np.linalg.norm(np.cross(axis, dt)), 1.0
            else:
                lin, ang = 1.0, 0.0
            radius = radii[j] + float(dt @ dt)
            length = max(length, max(lin, radius * ang))
        if min_len > length > 0:
            min_len = length
        w[i - 1] = length
        for k in range(i):
            if w[k] == 0:
                w[k] = min_len
    return [float(x) for x in w]


def main():
    robot = ur5()
    robot["weights"] = weighed_distance_weights(robot)
    for nb in (0, 1, 8, 32):
        m = dict(robot)
        m.update(boxes(nb))
        m["name"] = "ur5_boxes%d" % nb
        with open(os.path.join(HERE, "ur5_boxes%d.json" % nb), "w") as f:
            json.dump(m, f, indent=1)
    print("wrote ur5_boxes{0,1,8,32}.json; weights =", robot["weights"])


if __name__ == "__main__":
    main()
