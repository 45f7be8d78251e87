import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as ge
pkg = ge.load()
ur5 = pkg.ModelDescription.load("ur5_boxes8.json")
m = pkg.Model(ur5)
nodes = pkg.uniform_configs(ur5, 1 << 20, seed=81)
qs = pkg.uniform_configs(ur5, 8192, seed=82)
for ratio in ("8", "16", "32", "48"):
    os.environ["HPPB200_KNN_TC_RATIO"] = ratio
    nn = pkg.NearestNeighbor(m, 0)
    nn.append(nodes)
    for k in (1, 10, 32):
        for nq in (1024, 8192):
            nn.query(qs[:nq], k)
            ts = []
            for _ in range(7):
                t0 = time.perf_counter(); nn.query(qs[:nq], k); ts.append(time.perf_counter() - t0)
            st = nn.tc_stats()
            print("ratio", ratio, "k", k, "queries", nq, "%.3f ms" % (sorted(ts)[3] * 1e3), "stages", st["stages"], "redone", st["redone"], "groups/query %.0f" % (st["candidates"] / nq), flush=True)
    del nn
cand = pkg.uniform_configs(ur5, 400, seed=5)
ws = pkg.Workspace(m, 0)
free = cand[ws.validate_configs(cand, check_bounds=True)[0] == 1]
del os.environ["HPPB200_KNN_TC_RATIO"]
for n in (100000, 400000):
    eb = pkg.EngineBackend(pkg, m, 0, 1, 1e-3)
    gpu = pkg.KPrmStar(eb, ur5, free[0], free[1], n, seed=21)
    t0 = time.perf_counter(); gpu.solve_batch(); tg = time.perf_counter() - t0
    print("kPRM* N", n, "k", gpu.k, "engine %.2f s" % tg, "validated edges", gpu.validated_edges, "edges", gpu.roadmap.num_edges(), {k: round(v, 3) for k, v in gpu.timings.items()}, flush=True)
