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


def _pkg():
    """the package module (host-side C++ helpers of libhppb200.so: kprm_link_candidates, roadmap_merge_edges)"""
    from . import kprm_link_candidates, roadmap_merge_edges  # noqa: F401
    import sys
    return sys.modules[__package__]


class EngineBackend:
    """hpp-core's interfaces for one model, served by the engine (Workspace + NearestNeighbor)."""

    def __init__(self, pkg, model, device=0, method=1, param=1e-3):
        self.ws = pkg.Workspace(model, device)
        self.nn = pkg.NearestNeighbor(model, device)
        self.method, self.param = method, param

    def append(self, q, cc):
        self.nn.append(np.asarray(q, np.float64).reshape(1, -1), np.asarray([cc], np.int32))

    def append_many(self, qs, ccs):
        self.nn.append(np.asarray(qs, np.float64), np.asarray(ccs, np.int32))

    def relabel(self, old, new):
        self.nn.relabel(old, new)

    def set_labels(self, cc):
        self.nn.set_labels(cc)

    def nearest(self, qs, cc):
        idx, dist = self.nn.query(qs, 1, cc)
        return idx[:, 0], dist[:, 0]

    def knearest(self, qs, k, with_dist=False):
        """Roadmap::nearestNodes(q, k) over the whole roadmap, nearest first (-1 padded)."""
        idx, dist = self.nn.query(qs, k, -1)
        return (idx, dist) if with_dist else idx

    def config_valid(self, qs):
        """ConfigValidations::validate (collision validation + joint bounds as in the default problem)."""
        return self.ws.validate_configs(qs, check_bounds=True)[0].astype(bool)

    def lengths(self, q1, q2):
        return self.ws.straight_lengths(q1, q2)

    def validate(self, q1, q2, L, reverse):
        r = self.ws.validate_edges(self.method, self.param, q1, q2, L, reverse=reverse)
        return r["valid"].astype(bool), r["t_last_valid"]

    def path_eval(self, q1, q2, L, t):
        return self.ws.path_eval(q1, q2, L, t)

    def validate_nodes(self, a, b):
        """path validation of the straight paths between roadmap nodes a[i] -> b[i], end configurations gathered on the
        device from the mirror (hppb200_validate_edges_indexed)"""
        r = self.ws.validate_edges_indexed(self.nn, self.method, self.param, a, b)
        return r["valid"].astype(bool)


class Roadmap:
    """Nodes in insertion order, undirected edges, connected-component labels (src/roadmap.cc:74-142,209-255)."""

    def __init__(self, backend, nq):
        self.b = backend
        # nodes and labels grow by doubling (a vstack per node is 48 MB of copying per insertion at 1 M nodes)
        self._nodes = np.zeros((1024, nq))
        self._cc = np.zeros(1024, np.int64)
        self.n = 0
        self._edges, self._edge_set, self._blocks, self._set_upto = [], set(), [], 0
        self.next_cc = 1 << 20  # labels of components created by add_nodes (planners use small labels themselves)

    @property
    def nodes(self):
        return self._nodes[:self.n]

    @property
    def cc(self):
        """connected-component label per node (numpy view)"""
        return self._cc[:self.n]

    @cc.setter
    def cc(self, labels):
        labels = np.asarray(labels, np.int64)
        assert len(labels) == self.n
        self._cc[:self.n] = labels

    def _reserve(self, extra):
        need = self.n + extra
        if need > len(self._cc):
            cap = max(need, 2 * len(self._cc))
            nodes, cc = np.zeros((cap, self._nodes.shape[1])), np.zeros(cap, np.int64)
            nodes[:self.n], cc[:self.n] = self._nodes[:self.n], self._cc[:self.n]
            self._nodes, self._cc = nodes, cc

    def extend(self, qs, ccs):
        """bulk append of nodes with given labels (pre-filled roadmaps), no duplicate check"""
        qs = np.asarray(qs, np.float64).reshape(-1, self._nodes.shape[1])
        self._reserve(len(qs))
        self._nodes[self.n:self.n + len(qs)] = qs
        self._cc[self.n:self.n + len(qs)] = np.asarray(ccs, np.int64)
        self.n += len(qs)
        self.b.append_many(qs, np.asarray(ccs, np.int32))

    @property
    def edges(self):
        """(a, b) in insertion order; bulk additions are kept as arrays until somebody looks"""
        for blk in self._blocks:
            self._edges.extend(map(tuple, blk.tolist()))
        self._blocks = []
        return self._edges

    @property
    def edge_set(self):
        e = self.edges
        self._edge_set.update(e[self._set_upto:])
        self._set_upto = len(e)
        return self._edge_set

    def num_edges(self):
        return len(self._edges) + sum(len(b) for b in self._blocks)

    def add_node(self, q, cc):
        # Roadmap::addNode (:104-117): an existing node with exactly this configuration is reused
        if len(self.cc):
            i, _ = self.b.nearest(q.reshape(1, -1), -1)
            if np.array_equal(self.nodes[i[0]], q):
                return int(i[0])
        self._reserve(1)
        self._nodes[self.n] = q
        self._cc[self.n] = cc
        self.n += 1
        self.b.append(q, cc)
        return self.n - 1

    def add_nodes(self, qs):
        """Roadmap::addNode for a batch of NEW configurations (each its own connected component): one append; a
        configuration equal to an existing node is reused, as in :104-117 (looked up on the host here)."""
        seen = getattr(self, "_seen", None)
        if seen is None:
            seen = self._seen = {self.nodes[i].tobytes(): i for i in range(len(self.cc))}
        out, fresh = [], []
        for q in qs:
            key = q.tobytes()
            if key not in seen:
                seen[key] = len(self.cc) + len(fresh)
                fresh.append(q)
            out.append(seen[key])
        if fresh:
            ccs = list(range(self.next_cc, self.next_cc + len(fresh)))
            self.next_cc += len(fresh)
            self._reserve(len(fresh))
            self._nodes[self.n:self.n + len(fresh)] = np.asarray(fresh)
            self._cc[self.n:self.n + len(fresh)] = ccs
            self.n += len(fresh)
            self.b.append_many(np.asarray(fresh), ccs)
        return out

    def add_edges(self, pairs):
        """`add_edge` for many edges in the given order with the same final state (hppb200_roadmap_merge_edges: the
        component of an edge's second node joins the first node's, edge by edge); the device mirror receives the final
        labels in one call."""
        pairs = np.asarray(pairs, np.int64).reshape(-1, 2)
        if len(pairs) == 0:
            return
        self._blocks.append(pairs)
        new_cc = _pkg().roadmap_merge_edges(self.cc, pairs)      # union-find replay in insertion order (host C++)
        self.cc = new_cc
        self.b.set_labels(new_cc.astype(np.int32))

    def has_edge(self, a, b):
        return (a, b) in self.edge_set or (b, a) in self.edge_set

    def add_edge(self, a, b):
        self.edges.append((a, b))
        ca, cb = self.cc[a], self.cc[b]
        if ca != cb:  # ConnectedComponent::merge: the second component joins the first
            self.cc[self.cc == cb] = ca
            self.b.relabel(int(cb), int(ca))

    def add_node_and_edge(self, near, q):
        n = self.add_node(q, int(self.cc[near]))
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


class KPrmStar:
    """kPrmStar (reference src/path-planner/k-prm-star.cc:59-222) on the same backend interface.

    `solve_sequential` follows the reference state machine item by item: BUILD_ROADMAP (generateRandomConfig :100-132:
    shoot until valid, at most 10000 tries, addNode, until numberOfNodes nodes), LINK_NODES (linkNodes :134-151 +
    connectNodeToClosestNeighbors :153-193: every node in insertion order against its k = floor(2e ln N + .5) nearest
    nodes of the WHOLE roadmap, nearest first, skipping itself and neighbours that already have an edge to it;
    steer, validate forward, addEdges), CONNECT_INIT_GOAL (:195-217).
    `solve_batch` produces the same roadmap from four batches: configurations validated together (shots consumed in
    the same order), ONE k-NN call for all nodes, the edge validations the reference always performs, then the ones
    it performs only because the mirrored edge failed (a node pair (i, j), i > j, i among j's neighbours, is tried
    from i only if j -> i was invalid).  CONNECT_INIT_GOAL repeats validations whose (deterministic) outcome is known
    and cannot add an edge, so the batch form has nothing to do there."""

    K_PRM = 2.0 * np.e

    def __init__(self, backend, desc, q_init, q_goal, number_of_nodes, seed=1):
        if number_of_nodes <= 0:
            raise RuntimeError("kPrmStar: Number nodes should be positive, got %d" % number_of_nodes)  # :63-67
        self.b, self.desc, self.seed, self.n = backend, desc, seed, int(number_of_nodes)
        self.k = int(np.floor(self.K_PRM * np.log(float(self.n)) + 0.5))
        self.roadmap = Roadmap(backend, desc.nq)
        self.init = self.roadmap.add_node(np.asarray(q_init, np.float64), 0)
        self.goal = self.roadmap.add_node(np.asarray(q_goal, np.float64), 1)
        self.shots = 0
        self.symmetric_distance = int(desc.nq) == int(getattr(desc, "nv", desc.nq))  # no free-flyer joint

    def solved(self):
        return self.roadmap.cc[self.init] == self.roadmap.cc[self.goal]

    def _edge_valid(self, i, j):
        rm = self.roadmap
        qi, qj = rm.nodes[i].reshape(1, -1), rm.nodes[j].reshape(1, -1)
        return bool(self.b.validate(qi, qj, self.b.lengths(qi, qj), False)[0][0])

    def _connect(self, i, nbrs):
        rm = self.roadmap
        for j in nbrs:
            j = int(j)
            if j >= 0 and j != i and not rm.has_edge(j, i) and self._edge_valid(i, j):
                rm.add_edge(i, j)

    def solve_sequential(self):
        rm, b = self.roadmap, self.b
        while len(rm.cc) < self.n:
            tries, valid = 0, False
            while not valid and tries < 10000:
                q = uniform_configs(self.desc, 1, seed=self.seed, offset=self.shots)
                self.shots += 1
                valid = bool(b.config_valid(q)[0])
                tries += 1
            if not valid:
                raise RuntimeError("Failed to generate free configuration after 10000 trials.")
            rm.add_nodes(q)
        for i in range(len(rm.cc)):
            self._connect(i, b.knearest(rm.nodes[i].reshape(1, -1), self.k)[0])
        if not any(self.init in e for e in rm.edges):
            self._connect(self.init, b.knearest(rm.nodes[self.init].reshape(1, -1), self.k)[0])
        if not any(self.goal in e for e in rm.edges):
            self._connect(self.goal, b.knearest(rm.nodes[self.goal].reshape(1, -1), self.k)[0])
        return self.solved()

    def solve_batch(self, shoot_batch=4096):
        import time
        rm, b = self.roadmap, self.b
        tm = self.timings = {}
        t0 = time.perf_counter()
        invalid_run = 0
        while len(rm.cc) < self.n:
            q = uniform_configs(self.desc, shoot_batch, seed=self.seed, offset=self.shots)
            ok = b.config_valid(q)
            take = np.nonzero(ok)[0][:self.n - len(rm.cc)]
            used = shoot_batch if len(rm.cc) + len(take) < self.n else int(take[-1]) + 1  # shots the reference consumes
            run = invalid_run
            for v in ok[:used]:
                run = 0 if v else run + 1
                if run >= 10000:
                    raise RuntimeError("Failed to generate free configuration after 10000 trials.")
            invalid_run = run
            self.shots += used
            rm.add_nodes(q[take])
        n = len(rm.cc)
        tm["nodes (shoot + config validation + append)"] = time.perf_counter() - t0
        t0 = time.perf_counter()
        nbrs, dist = b.knearest(rm.nodes, self.k, with_dist=True)
        nbrs = np.asarray(nbrs, np.int64)
        tm["k-NN (one call)"] = time.perf_counter() - t0
        t0 = time.perf_counter()
        # directed candidates (i, j), j among i's neighbours: `first` = always tried, `second` = tried by the reference only
        # if the mirrored edge j -> i (j < i, i among j's neighbours) was invalid; hppb200_kprm_link_candidates (host C++,
        # csrc/planner_link.cpp) classifies all N k of them and returns, for a `second`, the slot of its mirror
        kk = nbrs.shape[1]
        symmetric = self.symmetric_distance and not getattr(self.desc, "has_freeflyer", False)
        kind, mirror = _pkg().kprm_link_candidates(nbrs, dist if symmetric else None, symmetric)
        kind, mirror, j_of = kind.ravel(), mirror.ravel(), nbrs.ravel()

        def validate(sel):
            a, c = sel // kk, j_of[sel]
            if len(a) == 0:
                return np.zeros(0, bool)
            if hasattr(b, "validate_nodes"):  # both ends live in the device mirror: send indices, not configurations
                return np.asarray(b.validate_nodes(a, c), bool)
            qa, qc = rm.nodes[a], rm.nodes[c]
            return np.asarray(b.validate(qa, qc, b.lengths(qa, qc), False)[0], bool)

        ok = np.zeros(len(j_of), bool)
        f_idx = np.nonzero(kind == 1)[0]
        s_idx = np.nonzero(kind == 2)[0]
        tm["candidate pairs (host)"] = time.perf_counter() - t0
        t0 = time.perf_counter()
        ok[f_idx] = validate(f_idx)
        s_idx = s_idx[~ok[mirror[s_idx]]]                        # the mirrored edge was found invalid
        ok[s_idx] = validate(s_idx)
        self.validated_edges = int(len(f_idx) + len(s_idx))
        tm["edge validation (two batches, host gathers included)"] = time.perf_counter() - t0
        t0 = time.perf_counter()
        e_idx = np.nonzero(ok)[0]                                # node by node, nearest neighbour first: the reference order
        rm.add_edges(np.stack([e_idx // kk, j_of[e_idx]], axis=1))
        tm["roadmap edges + components (host)"] = time.perf_counter() - t0
        return self.solved()


class _Stepper:
    """Common pieces of the two planners below: shots, new components, valid parts."""

    def __init__(self, backend, desc, q_init, q_goal, seed):
        self.b, self.desc, self.seed = backend, desc, seed
        self.roadmap = Roadmap(backend, desc.nq)
        self.init = self.roadmap.add_node(np.asarray(q_init, np.float64), 0)
        self.goal = self.roadmap.add_node(np.asarray(q_goal, np.float64), 1)
        self.shots = 0
        self.validations = 0  # path validations asked of the backend (batched forms validate speculatively: >= sequential)

    def solved(self):
        return self.roadmap.cc[self.init] == self.roadmap.cc[self.goal]

    def shoot(self, n=1):
        q = uniform_configs(self.desc, n, seed=self.seed, offset=self.shots)
        self.shots += n
        return q

    def components(self):
        """labels of the connected components in ascending order (the reference iterates a std::set of shared
        pointers, i.e. in address order: any fixed order gives the same nodes and edges up to insertion order)"""
        return sorted(set(self.roadmap.cc.tolist()))

    def new_component_node(self, q):
        """Roadmap::addNode(q): an existing node with this configuration, or a new node in a new component"""
        rm = self.roadmap
        n = rm.add_node(q, rm.next_cc)
        if n == rm.n - 1 and rm.cc[n] == rm.next_cc:
            rm.next_cc += 1
        return n

    def validate(self, q1, q2):
        """(L, ok, t_last, end of the valid part) of the straight paths q1[i] -> q2[i], one batch"""
        q1, q2 = np.asarray(q1, np.float64).reshape(-1, self.desc.nq), np.asarray(q2, np.float64).reshape(-1, self.desc.nq)
        L = self.b.lengths(q1, q2)
        ok, tl = self.b.validate(q1, q2, L, False)
        self.validations += len(L)
        ends = q2.copy()
        bad = np.nonzero(~ok)[0]
        if len(bad):
            ends[bad] = self.b.path_eval(q1[bad], q2[bad], L[bad], tl[bad])
        return L, ok, tl, ends


class DiffusingPlanner(_Stepper):
    """DiffusingPlanner::oneStep (reference src/diffusing-planner.cc:180-318, no constraints, default parameters
    extensionStepLength = extensionStepRatio = -1): extend every connected component towards one random configuration,
    then try to connect the new nodes with one another and with the nearest nodes of the other components.
    `one_step` validates path by path in the reference's order; `one_step_batch` submits the extensions as one edge
    batch and ALL connection attempts as a second one, then replays the reference's loop on the results (an attempt
    the reference skips because an earlier one merged the two components is discarded): same roadmap."""

    def _extend_phase(self, q_rand, batch):
        rm = self.roadmap
        ccs = self.components()
        nears = [int(self.b.nearest(q_rand.reshape(1, -1), c)[0][0]) for c in ccs]  # Roadmap::nearestNode(q_rand, cc)
        res = None
        if batch:
            res = self.validate(rm.nodes[nears], np.repeat(q_rand.reshape(1, -1), len(nears), axis=0))
        new_nodes, delayed = [], []
        for k, near in enumerate(nears):
            L, ok, tl, end = (x[k] for x in res) if batch else (x[0] for x in self.validate(rm.nodes[near], q_rand))
            t_final = L if ok else tl
            if t_final == 0.0:  # validPath->timeRange().second == path->timeRange().first
                continue
            q_new = end
            if (not ok) or not any(np.array_equal(rm.nodes[n], q_new) for n in new_nodes):  # belongs(q_new, newNodes)
                new_nodes.append(rm.add_node_and_edge(near, q_new))
            else:
                delayed.append((near, q_new))
        for near, q_new in delayed:
            n = self.new_component_node(q_new)
            rm.add_edge(near, n)
        return new_nodes, nears

    def _connect(self, a, q_b, node_b, res):
        """one connection attempt of the reference's second phase, given its validation result"""
        rm = self.roadmap
        L, ok, tl, end = res
        if ok:
            rm.add_edge(a, node_b)
        elif tl > 0:  # validPath && validPath->length() > 0
            rm.add_node_and_edge(a, end)

    def one_step(self, batch=False):
        rm = self.roadmap
        q_rand = self.shoot()[0]
        new_nodes, nears = self._extend_phase(q_rand, batch)
        # attempts in the reference's order: (n1, later new nodes...), then (n1, nearest neighbours of other components)
        attempts = []
        for i, n1 in enumerate(new_nodes):
            attempts += [(n1, n2, False) for n2 in new_nodes[i + 1:]]
            attempts += [(n1, n2, True) for n2 in nears]
        results = None
        if batch and attempts:
            a = np.array([t[0] for t in attempts])
            c = np.array([t[1] for t in attempts])
            results = self.validate(rm.nodes[a].copy(), rm.nodes[c].copy())
        for k, (n1, n2, check_cc) in enumerate(attempts):
            if check_cc and rm.cc[n1] == rm.cc[n2]:
                continue
            if np.array_equal(rm.nodes[n1], rm.nodes[n2]):
                continue  # (the reference asserts q1 != q2)
            r = tuple(x[k] for x in results) if results is not None else tuple(x[0] for x in self.validate(rm.nodes[n1], rm.nodes[n2]))
            self._connect(n1, rm.nodes[n2], n2, r)

    def one_step_batch(self):
        self.one_step(batch=True)

    def solve(self, max_steps=1000, batch=False):
        steps = 0
        while not self.solved() and steps < max_steps:
            self.one_step(batch)
            steps += 1
        return steps


class VisibilityPrmPlanner(_Stepper):
    """VisibilityPrmPlanner::oneStep / visibleFromCC (reference src/visibility-prm-planner.cc:73-186): a random valid
    configuration is tested for visibility (a valid straight path) from the GUARD nodes of every connected component;
    seen from none: it becomes a guard; seen from several: it becomes a connection node linked to the nearest visible
    guard of each.  `one_step` validates guard by guard, component by component; `one_step_batch` validates the paths
    to ALL guards of ALL components as one batch (visibleFromCC is 'one q against every guard of a component': a
    ready-made batch).  Guards of a component are visited in node order; the shortest valid path wins (strict <)."""

    def __init__(self, backend, desc, q_init, q_goal, seed=1):
        super().__init__(backend, desc, q_init, q_goal, seed)
        self.guard = {self.init: True, self.goal: True}  # nodeStatus_

    def _valid_shot(self, batch):
        """shoot until the configuration validations accept (the shots consumed are those of the sequential loop)"""
        while True:
            if not batch:
                q = self.shoot()[0]
                if self.b.config_valid(q.reshape(1, -1))[0]:
                    return q
                continue
            block = uniform_configs(self.desc, 64, seed=self.seed, offset=self.shots)
            ok = self.b.config_valid(block)
            first = int(np.argmax(ok)) if ok.any() else -1
            self.shots += 64 if first < 0 else first + 1
            if first >= 0:
                return block[first]

    def one_step(self, batch=False):
        rm = self.roadmap
        q = self._valid_shot(batch)
        guards = [n for n in range(rm.n) if self.guard.get(n, False)]
        ccs = self.components()
        by_cc = {c: [n for n in guards if rm.cc[n] == c] for c in ccs}
        res = None
        if batch and guards:
            order = [n for c in ccs for n in by_cc[c]]
            L, ok, _, _ = self.validate(np.repeat(q.reshape(1, -1), len(order), axis=0), rm.nodes[order])
            res = {n: (L[k], ok[k]) for k, n in enumerate(order)}
        delayed = []
        for c in ccs:  # visibleFromCC
            best, best_len = -1, np.inf
            for n in by_cc[c]:
                if res is not None:
                    L, ok = res[n]
                else:
                    Lv, okv, _, _ = self.validate(q, rm.nodes[n])
                    L, ok = Lv[0], okv[0]
                if ok and L < best_len:
                    best, best_len = n, L
            if best >= 0:
                delayed.append(best)
        if len(delayed) == 0:
            n = self.new_component_node(q)
            self.guard[n] = True
        if len(delayed) > 1:
            for near in delayed:
                n = self.new_component_node(q)
                self.guard[n] = False
                rm.add_edge(near, n)

    def one_step_batch(self):
        self.one_step(batch=True)

    def solve(self, max_steps=1000, batch=False):
        steps = 0
        while not self.solved() and steps < max_steps:
            self.one_step(batch)
            steps += 1
        return steps


class RandomShortcut:
    """pathOptimization::RandomShortcut::optimize (reference src/path-optimization/random-shortcut.cc:73-181) for a
    PathVector of StraightPaths given by its waypoints, no constraints: two random parameters t1 < t2, the three
    straight paths q0 -> q(t1) -> q(t2) -> q3 are validated, valid ones replace the corresponding parts, the result
    is kept if it is shorter; stop after NumberOfLoops iterations without improvement or when the last NumberOfLoops
    improvements gained less than 1e-4.
    `optimize(batch=False)` validates path by path like the reference.  `optimize(batch=True)` is the speculative form
    (SURVEY.md 8f-2): the loop ends after `projectionError` failures in a row, so that many iterations are drawn on the
    CURRENT path at once -- 2 evaluations + 3 validations each, one `path_eval` and one edge batch for all of them -- and
    replayed in order; the first improvement changes the path, the draws behind it are redone on the new path with the
    same random numbers.  Same result, `projectionError` times fewer validation calls."""

    def __init__(self, backend, desc, number_of_loops=5, seed=1):
        self.b, self.desc, self.n, self.seed = backend, desc, int(number_of_loops), seed
        self.draws = 0
        self.validation_calls = 0

    def _u(self, k):  # k-th pair of uniform numbers (the reference draws u2 first, then u1: shootTimes :183-195)
        from .model_io import splitmix64
        x = splitmix64(np.arange(2 * k, 2 * k + 2, dtype=np.uint64) + np.uint64(self.seed) * np.uint64(0x100000001B3))
        u = (x >> np.uint64(11)).astype(np.float64) * (2.0 ** -53)
        return u[0], u[1]

    def _seglens(self, W):
        return self.b.lengths(W[:-1], W[1:])

    @staticmethod
    def _rank(lens, t):
        """PathVector::rankAtParam (src/path-vector.cc:40-69): segment and local parameter of t"""
        r, local = 0, t
        while r + 1 < len(lens) and local > lens[r]:
            local -= lens[r]
            r += 1
        return r, min(local, lens[r])

    def _extract(self, W, lens, ta, qa, tb, qb):
        """waypoints of tmpPath->extract(ta, tb), ta <= tb: qa, the waypoints strictly inside, qb"""
        ra, _ = self._rank(lens, ta)
        rb, _ = self._rank(lens, tb)
        cum = np.concatenate([[0.0], np.cumsum(lens)])
        inner = [W[k] for k in range(ra + 1, rb + 1) if ta < cum[k] < tb]
        return [qa] + inner + [qb]

    def optimize(self, waypoints, batch=False):
        W = np.asarray(waypoints, np.float64)
        lens = self._seglens(W)
        length = [np.inf] * (self.n - 1) + [float(lens.sum())]
        perr, finished, result = self.n, False, None
        while not finished and perr != 0:
            depth = perr if batch else 1
            T3 = float(lens.sum())
            ts, idx = [], []
            for k in range(depth):
                u2, u1 = self._u(self.draws + k)
                a, c = T3 * u2, T3 * u1
                ts.append((min(a, c), max(a, c)))
            # q(t1), q(t2) of every draw: one path_eval over the segments they fall in
            seg, loc = [], []
            for t1, t2 in ts:
                for t in (t1, t2):
                    r, l = self._rank(lens, t)
                    seg.append(r)
                    loc.append(l)
            seg = np.array(seg)
            qs = self.b.path_eval(W[seg], W[seg + 1], lens[seg], np.array(loc)).reshape(depth, 2, -1)
            a = np.concatenate([np.stack([W[0], qs[k, 0], qs[k, 1]]) for k in range(depth)])
            c = np.concatenate([np.stack([qs[k, 0], qs[k, 1], W[-1]]) for k in range(depth)])
            L = self.b.lengths(a, c)
            ok, _ = self.b.validate(a, c, L, False)
            self.validation_calls += 1
            improved = False
            for k in range(depth):
                self.draws += 1
                t = [0.0, ts[k][0], ts[k][1], T3]
                q = [W[0], qs[k, 0], qs[k, 1], W[-1]]
                new = [q[0]]
                for i in range(3):
                    part = [q[i], q[i + 1]] if ok[3 * k + i] else self._extract(W, lens, t[i], q[i], t[i + 1], q[i + 1])
                    new += part[1:]
                newW = np.asarray(new)
                newlens = self._seglens(newW)
                newLength = float(newlens.sum())
                if length[-1] <= newLength:
                    perr -= 1
                    if perr == 0:
                        break
                    continue
                length = length[1:] + [newLength]
                finished = (length[0] - length[-1]) <= 1e-4 * length[-1]
                W, lens, result, perr, improved = newW, newlens, newW, self.n, True
                break  # the draws behind this one were made on the old path: redo them on the new one
            if not improved and not batch:
                continue
        return W if result is not None else np.asarray(waypoints, np.float64)


class SimpleShortcut:
    """pathOptimization::SimpleShortcut::optimize (reference src/path-optimization/simple-shortcut.cc:51-83) for a
    PathVector of StraightPaths given by its waypoints: a roadmap of the waypoints with the path's own edges
    (i -> i + 1), every pair (i, j >= i + 2) steered and validated, the valid ones added as edges i -> j, and the
    solution extracted by A* (src/astar.hh:96-136) with edge cost = path length and heuristic = distance to the goal.
    The N (N - 1) / 2 - (N - 1) validations are independent of each other: ONE edge batch (the reference validates them
    one by one).  The A* replay keeps the reference's container semantics (open list re-sorted with a stable sort by
    estimated cost every iteration, out-edges in insertion order), so that cost ties resolve the same way."""

    def __init__(self, backend):
        self.b = backend
        self.validated = 0

    def optimize(self, waypoints):
        W = np.asarray(waypoints, np.float64)
        n = len(W)
        if n < 3:
            return W
        a, c = np.triu_indices(n, 2)
        L = self.b.lengths(W[a], W[c])
        ok, _ = self.b.validate(W[a], W[c], L, False)
        self.validated += len(a)
        own = self.b.lengths(W[:-1], W[1:])
        h = self.b.lengths(W, np.repeat(W[-1:], n, axis=0))
        out = [[(i + 1, float(own[i]))] if i + 1 < n else [] for i in range(n)]
        for e in np.nonzero(ok)[0]:
            out[int(a[e])].append((int(c[e]), float(L[e])))
        return W[self._astar(out, h, n - 1)]

    @staticmethod
    def _astar(out, h, goal):
        open_, closed, parent = [0], set(), {}
        g, f = {0: 0.0}, {0: 0.0}  # (std::map::operator[]: the init node's entries are value-initialised)
        while open_:
            open_.sort(key=lambda x: f[x])  # std::list::sort is stable
            cur = open_[0]
            if cur == goal:
                path = [cur]
                while path[-1] in parent:
                    path.append(parent[path[-1]])
                return path[::-1]
            open_.pop(0)
            closed.add(cur)
            for child, cost in out[cur]:
                if child in closed:
                    continue
                tmp = g[cur] + cost
                new = child not in open_
                if new or tmp < g[child]:
                    parent[child], g[child], f[child] = cur, tmp, tmp + float(h[child])
                    if new:
                        open_.append(child)
        raise RuntimeError("A* failed to find a solution to the goal.")


class PartialShortcut:
    """pathOptimization::PartialShortcut::optimize with its default parameters (onlyFullShortcut = true; reference
    src/path-optimization/partial-shortcut.cc:84-98, 100-151, 186-223) for a PathVector of StraightPaths given by its
    waypoints, no constraints: joint after joint, the joint's configuration is replaced along the WHOLE path by its
    interpolation between the path's first and last configurations (at the ratio of the arc length reached at each
    waypoint), the resulting chain of straight paths is validated, and kept if valid.
    `optimize(batch=False)` validates one chain per joint like the reference.  `optimize(batch=True)` is the speculative
    form: the candidates of ALL remaining joints are built on the current path and validated as one
    edge batch; the first valid one in joint order is adopted (the reference would have adopted exactly that one, the
    failed ones before it are final), the candidates behind it are rebuilt on the new path.  Same result; the number
    of validation calls drops from the number of joints to the number of adopted joints + 1.
    (The random phase, optimizeRandom :225-330, runs only when onlyFullShortcut is switched off and draws from C
    rand(); it is not restated.)"""

    def __init__(self, backend, desc):
        self.b, self.desc = backend, desc
        self.validation_calls = 0
        self.failed_joints = []
        seg, r = [], 0
        for t in desc.joint_type:
            sz = 7 if int(t) == 6 else 1
            seg.append((r, sz))
            r += sz
        self.segments = seg

    @staticmethod
    def _total(lens):  # PathVector::appendPath accumulates timeRange().second leaf by leaf
        t = 0.0
        for x in lens:
            t += float(x)
        return t

    def _candidate(self, W, lens, seg):
        """generatePath(opted, joint, 0, q0, t3, q3): waypoints of the new chain, or None (single leaf: rkAtP2 == rkAtP1)"""
        n = len(W) - 1
        if n < 2:
            return None
        t3, t, us = self._total(lens), 0.0, []
        for i in range(n - 1):
            t += float(lens[i])
            us.append(t / (t3 - 0.0))
        us = np.asarray(us)
        # interpolate(q0, q3, u) of the whole configuration, of which the joint's segment is taken (the Lie-group
        # interpolation is joint-wise): the straight path q0 -> q3 over the parameter range [0, 1] evaluated at u
        m = len(us)
        full = self.b.path_eval(np.repeat(W[:1], m, axis=0), np.repeat(W[-1:], m, axis=0), np.ones(m), us)
        new = W.copy()
        r, sz = seg
        new[1:n, r:r + sz] = full[:, r:r + sz]
        return new

    def optimize(self, waypoints, batch=False):
        W = np.asarray(waypoints, np.float64)
        lens = self.b.lengths(W[:-1], W[1:])
        self.failed_joints = []
        j, J = 0, len(self.segments)
        while j < J:
            todo = range(j, J) if batch else range(j, j + 1)
            cands = [self._candidate(W, lens, self.segments[x]) for x in todo]
            if cands[0] is None:  # a single straight path: nothing to do for any joint
                self.failed_joints += list(range(j, J))
                break
            a = np.concatenate([c[:-1] for c in cands])
            c2 = np.concatenate([c[1:] for c in cands])
            L = self.b.lengths(a, c2)
            ok, _ = self.b.validate(a, c2, L, False)
            self.validation_calls += 1
            m = len(W) - 1
            adopted = False
            for x, cand in zip(todo, cands):
                sl = slice((x - j) * m, (x - j + 1) * m)
                if bool(np.all(ok[sl])):
                    W, lens, adopted = cand, L[sl], True
                    j = x + 1
                    break
                self.failed_joints.append(x)
            if not adopted:
                j = todo[-1] + 1
        return W
