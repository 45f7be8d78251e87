#!/usr/bin/env python3
"""kPRM* end to end on the three batched interfaces (KPrmStar.solve_batch), stage times.   python tools/kprm_run.py N [N ...]"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as ge
from oracle import binding as ob
from oracle_backend import OracleBackend
pkg = ge.load()
ur5 = pkg.ModelDescription.load("ur5_boxes8.json")
m = pkg.Model(ur5)
o = ob.Oracle(ur5)
cand = pkg.uniform_configs(ur5, 400, seed=5)
free = cand[o.validate_configs(cand, check_bounds=True)[0] == 1]
for n in [int(a) for a in sys.argv[1:]] or [17000]:
    eb = pkg.EngineBackend(pkg, m, 0, 1, 1e-3)
    gpu = pkg.KPrmStar(eb, ur5, free[0], free[1], n, seed=21)
    t0 = time.perf_counter(); solved = gpu.solve_batch(); tg = time.perf_counter() - t0
    print("kPRM* N", n, "k", gpu.k, "engine %.2f s" % tg, "validated edges", gpu.validated_edges, "edges", gpu.roadmap.num_edges(), "solved", solved,
          {k: round(v, 3) for k, v in gpu.timings.items()}, flush=True)
    if n <= 20000:
        cpu = pkg.KPrmStar(OracleBackend(ob.Oracle(ur5), 1, 1e-3), ur5, free[0], free[1], n, seed=21)
        t0 = time.perf_counter(); cpu.solve_batch(); tc = time.perf_counter() - t0
        print("   oracle (all host threads) %.2f s" % tc, "same roadmap", gpu.roadmap.edges == cpu.roadmap.edges and np.array_equal(gpu.roadmap.nodes, cpu.roadmap.nodes)
              and np.array_equal(gpu.roadmap.cc, cpu.roadmap.cc), flush=True)
