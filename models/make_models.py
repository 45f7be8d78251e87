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


def ellipsoid_hull(n, a, b, c):
    """n points on an ellipsoid (Fibonacci lattice): every point is a vertex of the convex hull."""
    k = np.arange(n) + 0.5
    phi = np.arccos(1 - 2 * k / n)
    th = np.pi * (1 + 5 ** 0.5) * k
    return np.stack([a * np.cos(th) * np.sin(phi), b * np.sin(th) * np.sin(phi), c * np.cos(phi)], axis=1)


def humanoid(freeflyer=False):
    """Humanoid-like TREE: waist (2) + neck (2) + 2 arms (7 each) + 2 legs (6 each) = 30 revolute joints,
    capsule / sphere / box / convex-mesh bodies (BASELINE configs[3]); with `freeflyer` the pelvis hangs on a
    free-flyer root joint (R^3 x SO(3), nq = 37, nv = 36), otherwise it is fixed to the world.  Dimensions are
    made up (roughly 1.6 m tall); the point is the structure: joint parents that are not the previous joint,
    convex meshes through GJK, hundreds of body pairs."""
    J = []  # name, parent, type, xyz
    def add(name, parent, typ, xyz):
        J.append((name, parent, typ, xyz))
        return len(J)
    root = add("root_joint", 0, FREEFLYER, (0, 0, 0)) if freeflyer else 0
    waist_y = add("waist_yaw", root, REV_Z, (0, 0, 0.10))
    waist_p = add("waist_pitch", waist_y, REV_Y, (0, 0, 0.05))
    neck_y = add("neck_yaw", waist_p, REV_Z, (0, 0, 0.45))
    neck_p = add("neck_pitch", neck_y, REV_Y, (0, 0, 0.05))
    limbs = {}
    for side, sy in (("l", 1.0), ("r", -1.0)):
        a = add(side + "_shoulder_pitch", waist_p, REV_Y, (0, sy * 0.20, 0.38))
        a = add(side + "_shoulder_roll", a, REV_X, (0, sy * 0.03, 0))
        a = add(side + "_shoulder_yaw", a, REV_Z, (0, 0, -0.05))
        ua = a
        a = add(side + "_elbow", a, REV_Y, (0, 0, -0.25))
        a = add(side + "_wrist_yaw", a, REV_Z, (0, 0, -0.12))
        fa = a
        a = add(side + "_wrist_pitch", a, REV_Y, (0, 0, -0.12))
        a = add(side + "_wrist_roll", a, REV_X, (0, 0, -0.02))
        limbs[side + "arm"] = (ua, fa, a)
    for side, sy in (("l", 1.0), ("r", -1.0)):
        a = add(side + "_hip_yaw", root, REV_Z, (0, sy * 0.09, -0.05))
        a = add(side + "_hip_roll", a, REV_X, (0, 0, -0.03))
        a = add(side + "_hip_pitch", a, REV_Y, (0, 0, -0.03))
        th = a
        a = add(side + "_knee", a, REV_Y, (0, 0, -0.38))
        sh = a
        a = add(side + "_ankle_pitch", a, REV_Y, (0, 0, -0.38))
        a = add(side + "_ankle_roll", a, REV_X, (0, 0, -0.02))
        limbs[side + "leg"] = (th, sh, a)
    verts = []
    def mesh(n, a, b, c):
        begin = len(verts)
        verts.extend(ellipsoid_hull(n, a, b, c).tolist())
        return begin, n
    G = []  # name, joint, shape, params, placement, vbegin, vcount
    G.append(("pelvis", root, BOX, (0.10, 0.14, 0.07), se3(t=(0, 0, 0.0)), 0, 0))
    vb, vc = mesh(24, 0.11, 0.17, 0.22)
    G.append(("torso", waist_p, CONVEX, (0, 0, 0), se3(t=(0, 0, 0.22)), vb, vc))
    G.append(("head", neck_p, SPHERE, (0.10,), se3(t=(0, 0, 0.10)), 0, 0))
    for side in ("l", "r"):
        ua, fa, hand = limbs[side + "arm"]
        G.append((side + "_upper_arm", ua, CAPSULE, (0.045, 0.10), se3(t=(0, 0, -0.125)), 0, 0))
        G.append((side + "_forearm", fa, CAPSULE, (0.04, 0.09), se3(t=(0, 0, 0.0)), 0, 0))
        vb, vc = mesh(16, 0.035, 0.05, 0.08)
        G.append((side + "_hand", hand, CONVEX, (0, 0, 0), se3(t=(0, 0, -0.08)), vb, vc))
    for side in ("l", "r"):
        th, sh, foot = limbs[side + "leg"]
        G.append((side + "_thigh", th, CAPSULE, (0.07, 0.15), se3(t=(0, 0, -0.19)), 0, 0))
        G.append((side + "_shin", sh, CAPSULE, (0.055, 0.15), se3(t=(0, 0, -0.19)), 0, 0))
        G.append((side + "_foot", foot, BOX, (0.11, 0.05, 0.025), se3(t=(0.03, 0, -0.055)), 0, 0))
        vb, vc = mesh(12, 0.04, 0.045, 0.03)
        G.append((side + "_toe_cap", foot, CONVEX, (0, 0, 0), se3(t=(0.13, 0, -0.05)), vb, vc))
    name_to = {g[0]: i for i, g in enumerate(G)}
    def prs(a_names, b_names):
        return [[name_to[a], name_to[b]] for a in a_names for b in b_names if a != b]
    arms_l, arms_r = ["l_upper_arm", "l_forearm", "l_hand"], ["r_upper_arm", "r_forearm", "r_hand"]
    legs_l, legs_r = ["l_thigh", "l_shin", "l_foot", "l_toe_cap"], ["r_thigh", "r_shin", "r_foot", "r_toe_cap"]
    pairs = (prs(["l_forearm", "l_hand", "r_forearm", "r_hand"], ["torso", "head", "pelvis"]) +
             prs(arms_l[1:], arms_r) + prs(["l_upper_arm"], arms_r[1:]) +
             prs(["l_forearm", "l_hand"], legs_l[:2] + legs_r[:2]) + prs(["r_forearm", "r_hand"], legs_l[:2] + legs_r[:2]) +
             prs(legs_l, legs_r) + prs(["l_shin", "l_foot", "l_toe_cap", "r_shin", "r_foot", "r_toe_cap"], ["pelvis"]))
    lim = {REV_X: 0.8, REV_Y: 1.4, REV_Z: 1.2}
    q_lower, q_upper = [], []
    for j in J:
        if j[2] == FREEFLYER:  # translation box; quaternion ranks bounded by +-1.01 as hpp-pinocchio sets them
            q_lower += [-0.4, -0.4, -0.2] + [-1.01] * 4
            q_upper += [0.4, 0.4, 0.3] + [1.01] * 4
        else:
            q_lower.append(-lim[j[2]])
            q_upper.append(lim[j[2]])
    return {
        "name": "humanoid_freeflyer" if freeflyer else "humanoid_fixed_pelvis",
        "joint_names": [j[0] for j in J],
        "joint_parent": [j[1] for j in J],
        "joint_type": [j[2] for j in J],
        "joint_placement": [se3(None, j[3]) for j in J],
        "q_lower": q_lower,
        "q_upper": q_upper,
        "geom_names": [g[0] for g in G],
        "geom_joint": [g[1] for g in G],
        "geom_shape": [g[2] for g in G],
        "geom_params": [list(g[3]) + [0.0] * (4 - len(g[3])) for g in G],
        "geom_placement": [g[4] for g in G],
        "geom_vert_begin": [g[5] for g in G],
        "geom_vert_count": [g[6] for g in G],
        "self_pairs": pairs,
        "vertices": verts,
        "break_distance": 1e-2,
    }


def humanoid_obstacles(n, verts, seed=77):
    """n obstacles around the humanoid: boxes and (every third one) a convex mesh appended to `verts`."""
    rng = np.random.default_rng(seed)
    out = {"obst_names": [], "obst_placement": [], "obst_shape": [], "obst_params": [],
           "obst_vert_begin": [], "obst_vert_count": []}
    for k in range(n):
        rad, ang = rng.uniform(0.35, 0.9), rng.uniform(-np.pi, np.pi)
        c = (rad * np.cos(ang), rad * np.sin(ang), rng.uniform(-0.7, 0.9))
        out["obst_names"].append("obst_%d" % k)
        out["obst_placement"].append(se3(random_rotation(rng), c))
        if k % 3 == 2:
            begin = len(verts)
            verts.extend(ellipsoid_hull(20, *rng.uniform(0.08, 0.2, size=3)).tolist())
            out["obst_shape"].append(CONVEX)
            out["obst_params"].append([0.0] * 4)
            out["obst_vert_begin"].append(begin)
            out["obst_vert_count"].append(20)
        else:
            out["obst_shape"].append(BOX)
            out["obst_params"].append([float(x) / 2 for x in rng.uniform(0.1, 0.35, size=3)] + [0.0])
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
        elif s == CONVEX:
            vb, vc = model["geom_vert_begin"][g], model["geom_vert_count"][g]
            h = np.abs(np.array(model["vertices"][vb:vb + vc])).max(axis=0)
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
            if jt == FREEFLYER:  # rotation sub-Jacobian of R^3 x SO(3): singular values of [dt]x and of R
                lin, ang = np.linalg.norm(dt), 1.0
            elif jt <= REV_Z:
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
    for nb in (6, 16):
        h = humanoid()
        h["vertices"] = list(h["vertices"])
        h.update(humanoid_obstacles(nb, h["vertices"]))
        h["weights"] = weighed_distance_weights(h)
        h["name"] = "humanoid_obst%d" % nb
        with open(os.path.join(HERE, "humanoid_obst%d.json" % nb), "w") as f:
            json.dump(h, f, indent=1)
    print("wrote humanoid_obst{6,16}.json: %d joints, %d geometries, %d self pairs" % (
        len(h["joint_parent"]), len(h["geom_joint"]), len(h["self_pairs"])))
    h = humanoid(freeflyer=True)
    h["vertices"] = list(h["vertices"])
    h.update(humanoid_obstacles(6, h["vertices"]))
    h["weights"] = weighed_distance_weights(h)
    h["name"] = "humanoid_ff_obst6"
    with open(os.path.join(HERE, "humanoid_ff_obst6.json"), "w") as f:
        json.dump(h, f, indent=1)
    print("wrote humanoid_ff_obst6.json: %d joints (free-flyer root), nq = %d" % (len(h["joint_parent"]), len(h["q_lower"])))


if __name__ == "__main__":
    main()
