"""One warm-up + two timed tensor-core k-NN queries (k = 10, k = 1) over a 1 M-node mirror: the ncu target."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as ge
pkg = ge.load()
ur5 = pkg.ModelDescription.load("ur5_boxes8.json")
m = pkg.Model(ur5)
nodes = pkg.uniform_configs(ur5, 1 << 20, seed=81)
qs = pkg.uniform_configs(ur5, 1024, seed=82)
nn = pkg.NearestNeighbor(m, 0)
nn.append(nodes)
for k in (10, 1, 10, 1):
    t0 = time.perf_counter()
    nn.query(qs, k)
    print("k", k, "%.3f ms" % ((time.perf_counter() - t0) * 1e3), nn.tc_stats(), flush=True)
