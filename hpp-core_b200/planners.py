"""Host-side planner front-end over the three accelerated interfaces.

`BiRRTPlanner.one_step` restates the control flow of the reference's BiRRTPlanner::oneStep
(src/bi-rrt-planner.cc:109-170, no constraints: extendInternal :76-97 is the straight steering method) on top of
a backend that provides exactly what hpp-core asks of ConfigValidation / PathValidation / NearestNeighbor /
steeringMethod::Straight: nearest node of a connected component, straight-path length, path validation with
valid part.  No arithmetic happens here; `EngineBackend` forwards to the C ABI.  `one_step_batch` is the batch
extension: the independent (shoot, nearest, steer, validate) chains of several steps are submitted as ONE
k-NN query batch and ONE edge batch per phase.

The roadmap (nodes, edges, connected components) stays on the host, as in the reference.
"""
import numpy as np

from .model_io import uniform_configs


class EngineBackend:
    """hpp-core's interfaces for one model, served by the engine (Workspace + NearestNeighbor)."""

    def __init__(self, pkg, model, device=0, method=1, param=1e-3):
        self.ws = pkg.Workspace(model, device)
        self.nn = pkg.NearestNeighbor(model, device)
        self.method, self.param = method, param

    def append(self, q, cc):
        self.nn.append(np.asarray(q, np.float64).reshape(1, -1), np.asarray([cc], np.int32))

    def relabel(self, old, new):
        self.nn.relabel(old, new)

    def nearest(self, qs, cc):
        idx, dist = self.nn.query(qs, 1, cc)
        return idx[:, 0], dist[:, 0]

    def lengths(self, q1, q2):
        return self.ws.straight_lengths(q1, q2)

    def validate(self, q1, q2, L, reverse):
        r = self.ws.validate_edges(self.method, self.param, q1, q2, L, reverse=reverse)
        return r["valid"].astype(bool), r["t_last_valid"]

    def path_eval(self, q1, q2, L, t):
        return self.ws.path_eval(q1, q2, L, t)


class Roadmap:
    """Nodes in insertion order, undirected edges, connected-component labels (src/roadmap.cc:74-142,209-255)."""

    def __init__(self, backend, nq):
        self.b = backend
        self.nodes = np.zeros((0, nq))
        self.cc = []
        self.edges = []

    def add_node(self, q, cc):
        # Roadmap::addNode (:104-117): an existing node with exactly this configuration is reused
        if len(self.cc):
            i, _ = self.b.nearest(q.reshape(1, -1), -1)
            if np.array_equal(self.nodes[i[0]], q):
                return int(i[0])
        self.nodes = np.vstack([self.nodes, q.reshape(1, -1)])
        self.cc.append(cc)
        self.b.append(q, cc)
        return len(self.cc) - 1

    def add_edge(self, a, b):
        self.edges.append((a, b))
        ca, cb = self.cc[a], self.cc[b]
        if ca != cb:  # ConnectedComponent::merge: the second component joins the first
            self.cc = [ca if c == cb else c for c in self.cc]
            self.b.relabel(cb, ca)

    def add_node_and_edge(self, near, q):
        n = self.add_node(q, self.cc[near])
        self.add_edge(near, n)
        return n


class BiRRTPlanner:
    def __init__(self, backend, desc, q_init, q_goal, seed=1):
        self.b, self.desc, self.seed = backend, desc, seed
        self.roadmap = Roadmap(backend, desc.nq)
        self.init = self.roadmap.add_node(np.asarray(q_init, np.float64), 0)
        self.goal = self.roadmap.add_node(np.asarray(q_goal, np.float64), 1)
        self.shots = 0

    def solved(self):
        return self.roadmap.cc[self.init] == self.roadmap.cc[self.goal]

    def shoot(self, n=1):
        q = uniform_configs(self.desc, n, seed=self.seed, offset=self.shots)
        self.shots += n
        return q

    def _valid_part(self, q1, q2, L, ok, t_last, reverse):
        """(initial, end, length) of validPart: the path itself, or extract(0, t_last) / extract(t_last, L)."""
        if ok:
            return q1, q2, L
        t0, t1 = (t_last, L) if reverse else (0.0, t_last)
        q = self.b.path_eval(np.stack([q1, q1]), np.stack([q2, q2]), np.array([L, L]), np.array([t0, t1]))
        return q[0], q[1], abs(t1 - t0)

    def one_step(self):
        """BiRRTPlanner::oneStep (src/bi-rrt-planner.cc:109-170)."""
        rm, b = self.roadmap, self.b
        q_rand = self.shoot()[0]
        start_cc, goal_cc = rm.cc[self.init], rm.cc[self.goal]
        near = int(b.nearest(q_rand.reshape(1, -1), start_cc)[0][0])
        q_near = rm.nodes[near]
        L = float(b.lengths(q_near.reshape(1, -1), q_rand.reshape(1, -1))[0])
        ok, tl = b.validate(q_near.reshape(1, -1), q_rand.reshape(1, -1), np.array([L]), False)
        valid_from_start = bool(ok[0])
        _, q_new, length = self._valid_part(q_near, q_rand, L, valid_from_start, float(tl[0]), False)
        connected, reached = False, -1
        if length != 0.0:
            connected = True
            reached = rm.add_node_and_edge(near, q_new)
        # the end component
        near = int(b.nearest(q_rand.reshape(1, -1), goal_cc)[0][0])
        q_near = rm.nodes[near]
        L = float(b.lengths(q_rand.reshape(1, -1), q_near.reshape(1, -1))[0])
        ok, tl = b.validate(q_rand.reshape(1, -1), q_near.reshape(1, -1), np.array([L]), True)
        if ok[0] and valid_from_start and reached >= 0:
            rm.add_edge(reached, near)
            return
        q_new_end, _, length = self._valid_part(q_rand, q_near, L, bool(ok[0]), float(tl[0]), True)
        if length != 0.0:
            new_node = rm.add_node_and_edge(near, q_new_end)
            if connected:
                L = float(b.lengths(q_new.reshape(1, -1), q_new_end.reshape(1, -1))[0])
                ok, _ = b.validate(q_new.reshape(1, -1), q_new_end.reshape(1, -1), np.array([L]), False)
                if ok[0]:
                    rm.add_edge(reached, new_node)

    def one_step_batch(self, nshots):
        """`nshots` extensions towards independent random configurations per tree, each phase one batch:
        nearest nodes (one k-NN call per tree), straight lengths, path validation (one call per direction),
        valid parts (one path_eval call), then the connection attempts of the new node pairs."""
        rm, b = self.roadmap, self.b
        q_rand = self.shoot(nshots)
        start_cc, goal_cc = rm.cc[self.init], rm.cc[self.goal]
        ns, _ = b.nearest(q_rand, start_cc)
        ng, _ = b.nearest(q_rand, goal_cc)
        qs, qg = rm.nodes[ns], rm.nodes[ng]
        Ls, Lg = b.lengths(qs, q_rand), b.lengths(q_rand, qg)
        oks, tls = b.validate(qs, q_rand, Ls, False)
        okg, tlg = b.validate(q_rand, qg, Lg, True)
        ends = b.path_eval(qs, q_rand, Ls, np.where(oks, Ls, tls))        # validPart.end() of the start side
        begins = b.path_eval(q_rand, qg, Lg, np.where(okg, 0.0, tlg))     # validPart.initial() of the end side
        new_s, new_g = [], []
        for i in range(nshots):
            a = rm.add_node_and_edge(int(ns[i]), ends[i]) if (oks[i] and Ls[i] != 0) or (not oks[i] and tls[i] != 0) else -1
            g = rm.add_node_and_edge(int(ng[i]), begins[i]) if (okg[i] and Lg[i] != 0) or (not okg[i] and tlg[i] != Lg[i]) else -1
            new_s.append(a)
            new_g.append(g)
        pairs = [(a, g) for a, g in zip(new_s, new_g) if a >= 0 and g >= 0]
        if pairs and not self.solved():
            a = np.array([p[0] for p in pairs])
            g = np.array([p[1] for p in pairs])
            L = b.lengths(rm.nodes[a], rm.nodes[g])
            ok, _ = b.validate(rm.nodes[a], rm.nodes[g], L, False)
            for k in np.nonzero(ok)[0]:
                rm.add_edge(int(a[k]), int(g[k]))

    def solve(self, max_steps=10000, batch=0):
        steps = 0
        while not self.solved() and steps < max_steps:
            if batch:
                self.one_step_batch(batch)
            else:
                self.one_step()
            steps += 1
        return steps

    def path(self):
        """Node indices from init to goal over the roadmap edges (breadth first)."""
        adj = {}
        for a, c in self.roadmap.edges:
            adj.setdefault(a, []).append(c)
            adj.setdefault(c, []).append(a)
        prev, todo = {self.init: -1}, [self.init]
        while todo:
            n = todo.pop(0)
            if n == self.goal:
                break
            for m in adj.get(n, []):
                if m not in prev:
                    prev[m] = n
                    todo.append(m)
        if self.goal not in prev:
            return []
        out, n = [], self.goal
        while n >= 0:
            out.append(n)
            n = prev[n]
        return out[::-1]
