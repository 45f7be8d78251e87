import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
pkg = ge.load()
from oracle import binding as ob
sys.path.insert(0, os.path.join(ROOT, "tests"))
desc = pkg.ModelDescription.load("ur5_boxes8.json")
o = ob.Oracle(desc)
m = pkg.Model(desc)
golden = np.load(os.path.join(ROOT, "tests", "golden", "ur5_random_numbers_m1.npy"))
def edges(n, seed, short=False):
    q = pkg.uniform_configs(desc, 2 * n, seed=seed)
    q1, q2 = q[0::2].copy(), q[1::2].copy()
    if short:
        d = q2 - q1
        q2 = np.clip(q1 + 0.3 * d / np.linalg.norm(d, axis=1, keepdims=True), -np.pi, np.pi)
    return q1, q2
g1, g2 = golden[0:200:2], golden[1:200:2]
a1, a2 = edges(3000, 33)
b1, b2 = edges(2000, 34, short=True)
z = pkg.uniform_configs(desc, 300, seed=35)
q1, q2 = np.concatenate([g1, a1, b1, z]), np.concatenate([g2, a2, b2, z])
L = o.straight_lengths(q1, q2)
os.environ["HPPB200_WF_MIN"] = "1"
for rev in (False, True):
    e = o.validate_edges(1, 1e-3, q1, q2, L, reverse=rev)
    for form in ("1", "2"):
        os.environ["HPPB200_WF_CONTROL"] = form
        ws = pkg.Workspace(m, 0)
        r = ws.validate_edges(1, 1e-3, q1, q2, L, reverse=rev)
        bad = np.nonzero(~((r["t_last_valid"] == e["t_last_valid"]) & (r["valid"] == e["valid"]) & (r["pair"] == e["pair"])))[0]
        print("reverse", rev, "form", form, "bad", len(bad), bad[:10], ws.stats())
        for i in bad[:6]:
            print("  edge", i, "L", L[i], "gpu", r["valid"][i], r["t_fail"][i], r["t_last_valid"][i], r["pair"][i], r["reason"][i],
                  "oracle", e["valid"][i], e["t_fail"][i], e["t_last_valid"][i], e["pair"][i], e["reason"][i], e["samples"][i])
