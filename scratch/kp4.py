import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as ge
pkg = ge.load()
ur5 = pkg.ModelDescription.load("ur5_boxes8.json")
m = pkg.Model(ur5)
cand = pkg.uniform_configs(ur5, 400, seed=5)
ws = pkg.Workspace(m, 0)
free = cand[ws.validate_configs(cand, check_bounds=True)[0] == 1]
for n in (17000, 100000, 400000, 1000000):
    eb = pkg.EngineBackend(pkg, m, 0, 1, 1e-3)
    gpu = pkg.KPrmStar(eb, ur5, free[0], free[1], n, seed=21)
    t0 = time.perf_counter(); gpu.solve_batch(shoot_batch=1 << 18); tg = time.perf_counter() - t0
    print("kPRM* N", n, "k", gpu.k, "engine %.2f s" % tg, "validated edges", gpu.validated_edges, "edges", gpu.roadmap.num_edges(), "solved", gpu.solved(), {k: round(v, 3) for k, v in gpu.timings.items()}, eb.nn.tc_stats(), flush=True)
    del gpu, eb
