"""The planner backend interface of hpp-core_b200/planners.py answered by the CPU oracle (test infrastructure:
only tests may import oracle/)."""
import numpy as np


def _threads(n):
    """item-by-item planner calls stay on one thread (an OpenMP team per single edge only spins); batches use all"""
    return 1 if n < 64 else 0


class OracleBackend:
    """The same calls as EngineBackend, answered by the CPU oracle."""

    def __init__(self, o, method, param):
        self.o, self.method, self.param = o, method, param

    def _track(self, ccs):
        self._labels = np.concatenate([getattr(self, "_labels", np.zeros(0, np.int64)), np.asarray(ccs, np.int64).ravel()])

    def append(self, q, cc):
        self._track([cc])
        self.o.knn_append(np.asarray(q, np.float64).reshape(1, -1), np.asarray([cc], np.int32))

    def append_many(self, qs, ccs):
        self._track(ccs)
        self.o.knn_append(np.asarray(qs, np.float64), np.asarray(ccs, np.int32))

    def relabel(self, old, new):
        self._labels[self._labels == old] = new
        self.o.knn_relabel(old, new)

    def set_labels(self, cc):
        """final labels of a bulk merge, replayed as (old -> new) relabels of the nodes whose label changed"""
        cc = np.asarray(cc, np.int64)
        old = getattr(self, "_labels", None)
        if old is None or len(old) != len(cc):
            raise RuntimeError("set_labels needs the labels given at append time")
        for o_, n_ in sorted({(int(a), int(b)) for a, b in zip(old, cc) if a != b}):
            self.o.knn_relabel(o_, n_)
        self._labels = cc.copy()

    def nearest(self, qs, cc):
        idx, dist = self.o.knn_query(qs, 1, cc, nthreads=_threads(len(qs)))
        return idx[:, 0], dist[:, 0]

    def knearest(self, qs, k, with_dist=False):
        idx, dist = self.o.knn_query(np.asarray(qs, np.float64), k, -1, nthreads=_threads(len(qs)))
        srt = np.lexsort((idx, dist), axis=1)  # ties of std::priority_queue canonicalised by node index
        idx, dist = np.take_along_axis(idx, srt, axis=1), np.take_along_axis(dist, srt, axis=1)
        return (idx, dist) if with_dist else idx

    def config_valid(self, qs):
        return self.o.validate_configs(qs, check_bounds=True, nthreads=_threads(len(qs)))[0].astype(bool)

    def lengths(self, q1, q2):
        return self.o.straight_lengths(q1, q2)

    def validate(self, q1, q2, L, reverse):
        r = self.o.validate_edges(self.method, self.param, q1, q2, L, reverse=reverse, nthreads=_threads(len(L)))
        return r["valid"].astype(bool), r["t_last_valid"]

    def path_eval(self, q1, q2, L, t):
        return self.o.interpolate(q1, q2, L, t)
