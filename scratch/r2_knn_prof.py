import os, sys, numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import __graft_entry__ as ge
pkg = ge.load()
desc = pkg.ModelDescription.load("ur5_boxes8.json"); model = pkg.Model(desc)
nn = pkg.NearestNeighbor(model, 0)
nn.append(pkg.uniform_configs(desc, 1 << 20, seed=12347))
qs = pkg.uniform_configs(desc, 1024, seed=12348)
for _ in range(3):
    nn.query(qs, 10)
