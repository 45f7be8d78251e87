import sys, os, time
sys.path.insert(0, "/root/repo")
import numpy as np
import __graft_entry__ as ge
pkg = ge.load()
desc = pkg.ModelDescription.load("ur5_boxes8.json"); model = pkg.Model(desc)
os.environ["HPPB200_KNN_TIMING"] = "1"
nn = pkg.NearestNeighbor(model, 0); nn.append(pkg.uniform_configs(desc, 1 << 20, seed=3))
q = pkg.uniform_configs(desc, 1, seed=5)
for k in (1, 10):
    for _ in range(30): nn.query(q, k)
    t0 = time.perf_counter()
    for _ in range(300): nn.query(q, k)
    dt = (time.perf_counter() - t0) / 300
    print("ctas", os.environ.get("HPPB200_KNN_LAT_CTAS", "4"), "k", k, "call %.1f us" % (dt * 1e6), "kernel", nn.latency_time() if hasattr(nn, "latency_time") else "")
