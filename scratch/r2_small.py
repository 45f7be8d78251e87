"""Small-batch Progressive latency: CTA-per-edge kernel vs the CTA engine, parity vs the oracle."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
pkg = ge.load()
from oracle import binding as ob
desc = pkg.ModelDescription.load(sys.argv[1] if len(sys.argv) > 1 else "ur5_boxes8.json")
o = ob.Oracle(desc)
m = pkg.Model(desc)
for n in (1, 64, 512, 4096, 32768):
    q = pkg.uniform_configs(desc, 2 * n, seed=77)
    q1, q2 = q[0::2].copy(), q[1::2].copy()
    L = o.straight_lengths(q1, q2)
    e = o.validate_edges(1, 1e-3, q1, q2, L) if n <= 4096 else None
    for tm in ("0", "100000000"):
        os.environ["HPPB200_PROG_TAIL_MAX"] = tm
        ws = pkg.Workspace(m, 0)
        best = 1e9
        for it in range(4):
            t0 = time.perf_counter(); r = ws.validate_edges(1, 1e-3, q1, q2, L); best = min(best, time.perf_counter() - t0)
        ok = "-" if e is None else all(np.array_equal(r[k], e[k], equal_nan=True) for k in ("valid", "reason", "t_fail", "t_last_valid", "pair"))
        rr = ws.validate_edges(1, 1e-3, q1, q2, L, reverse=True)
        ok2 = "-" if e is None else all(np.array_equal(rr[k], ee, equal_nan=True) for k, ee in o.validate_edges(1, 1e-3, q1, q2, L, reverse=True).items() if k in ("valid", "reason", "t_fail", "t_last_valid", "pair"))
        print("n=%6d tail_max=%-9s best %.3f ms (%.1f us/edge) parity fwd %s rev %s" % (n, tm, best * 1e3, best * 1e6 / n, ok, ok2), flush=True)
