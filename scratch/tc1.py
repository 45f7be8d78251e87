"""Bring-up of the tensor-core k-NN candidate pass: accumulator tile vs numpy, parity vs the exact scan, timing."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as ge
pkg = ge.load()
ur5 = pkg.ModelDescription.load("ur5_boxes8.json")
m = pkg.Model(ur5)
w = np.asarray(ur5.weights)
N = int(os.environ.get("N", 200000))
nodes = pkg.uniform_configs(ur5, N, seed=81)
qs = pkg.uniform_configs(ur5, 1024, seed=82)
nn = pkg.NearestNeighbor(m, 0)
nn.append(nodes)
ref = (((qs[:128, None, :] - nodes[None, :256, :]) * w) ** 2).sum(-1)
for swap in (0, 1):
    t = nn.tc_debug_tile(qs[:128], swap)
    err = np.abs(t - ref)
    print("swap", swap, "max abs err", err.max(), "median", np.median(err), "tile[0,:4]", t[0, :4], "ref", ref[0, :4], flush=True)
os.environ["HPPB200_KNN_TC"] = "0"
ex = pkg.NearestNeighbor(m, 0)
ex.append(nodes)
del os.environ["HPPB200_KNN_TC"]
for k in (1, 10, 32):
    for nqry in (64, 200, 1024):
        gi, gd = nn.query(qs[:nqry], k)
        st = nn.tc_stats()
        ei, ed = ex.query(qs[:nqry], k)
        print("k", k, "queries", nqry, "idx equal", np.array_equal(gi, ei), "dist equal", np.array_equal(gd, ed), st, ex.tc_stats()["used"], flush=True)
for k in (1, 10):
    for obj, name in ((nn, "tc"), (ex, "exact")):
        obj.query(qs, k)
        t0 = time.perf_counter()
        for _ in range(5):
            obj.query(qs, k)
        dt = (time.perf_counter() - t0) / 5
        print(name, "k", k, "1024 queries over", N, "nodes: %.3f ms  %.0f queries/s" % (dt * 1e3, 1024 / dt), flush=True)
