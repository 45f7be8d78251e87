"""k > 32 on the tensor-core path (timing + parity vs exact scan) and kPRM* end to end: engine vs oracle."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import __graft_entry__ as ge
from oracle import binding as ob
from oracle_backend import OracleBackend
pkg = ge.load()
ur5 = pkg.ModelDescription.load("ur5_boxes8.json")
m = pkg.Model(ur5)
nodes = pkg.uniform_configs(ur5, 1 << 20, seed=81)
qs = pkg.uniform_configs(ur5, 1024, seed=82)
nn = pkg.NearestNeighbor(m, 0)
nn.append(nodes)
os.environ["HPPB200_KNN_TC"] = "0"
ex = pkg.NearestNeighbor(m, 0)
ex.append(nodes)
del os.environ["HPPB200_KNN_TC"]
for k in (33, 75, 128):
    gi, gd = nn.query(qs, k)
    st = nn.tc_stats()
    t0 = time.perf_counter(); nn.query(qs, k); t1 = time.perf_counter() - t0
    nq_ex = 64
    t0 = time.perf_counter(); ei, ed = ex.query(qs[:nq_ex], k); t2 = time.perf_counter() - t0
    print("k", k, "tc 1024 queries %.3f ms (%.0f q/s)" % (t1 * 1e3, 1024 / t1), st, "exact scan %d queries %.1f ms (%.0f q/s)" % (nq_ex, t2 * 1e3, nq_ex / t2),
          "equal", np.array_equal(gi[:nq_ex], ei), np.array_equal(gd[:nq_ex], ed), flush=True)
o = ob.Oracle(ur5)
cand = pkg.uniform_configs(ur5, 400, seed=5)
free = cand[o.validate_configs(cand, check_bounds=True)[0] == 1]
for n in (17000, 100000, 400000):
    eb = pkg.EngineBackend(pkg, m, 0, 1, 1e-3)
    gpu = pkg.KPrmStar(eb, ur5, free[0], free[1], n, seed=21)
    t0 = time.perf_counter(); gpu.solve_batch(); tg = time.perf_counter() - t0
    print("kPRM* N", n, "k", gpu.k, "engine %.2f s" % tg, "validated edges", gpu.validated_edges, "edges", len(gpu.roadmap.edges), eb.nn.tc_stats(), flush=True)
    if n <= 20000:
        cpu = pkg.KPrmStar(OracleBackend(ob.Oracle(ur5), 1, 1e-3), ur5, free[0], free[1], n, seed=21)
        t0 = time.perf_counter(); cpu.solve_batch(); tc = time.perf_counter() - t0
        print("   oracle (all host threads) %.2f s" % tc, "same roadmap", gpu.roadmap.edges == cpu.roadmap.edges and np.array_equal(gpu.roadmap.nodes, cpu.roadmap.nodes), flush=True)
