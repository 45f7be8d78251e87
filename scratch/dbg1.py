import sys, numpy as np
sys.path.insert(0, '/root/repo')
import __graft_entry__ as ge
pkg = ge.load()
from oracle import binding as ob
desc = pkg.ModelDescription.load("ur5_boxes8.json")
m = pkg.Model(desc); ws = pkg.Workspace(m, 0); o = ob.Oracle(desc)
q = pkg.uniform_configs(desc, 8000, seed=31)
q1, q2 = q[0::2].copy(), q[1::2].copy()
L = o.straight_lengths(q1, q2)
for rev in (False, True):
    r = ws.validate_edges(2, 0.0, q1, q2, L, reverse=rev)
    e = o.validate_edges(2, 0.0, q1, q2, L, reverse=rev)
    bad = np.nonzero(r["valid"] != e["valid"])[0]
    print("rev", rev, "nbad", len(bad), "reasons gpu", np.bincount(r["reason"], minlength=4), "oracle", np.bincount(e["reason"], minlength=4))
    for i in bad[:10]:
        print(i, "gpu", r["valid"][i], r["reason"][i], r["t_fail"][i], r["t_last_valid"][i], "| oracle", e["valid"][i], e["reason"][i], e["t_fail"][i], e["t_last_valid"][i], "samples", e["samples"][i], "L", L[i])
    print("oracle samples max", e["samples"].max())
