#!/usr/bin/env python3
"""Headline benchmark: validated configs/sec (+ edges/sec) on UR5 + 8 box obstacles.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Workload at N=1 (BASELINE.json configs[1]): 1 M uniform random UR5 configurations through
CollisionValidation (FK + narrow phase) per step, synthetic inputs from the counter-based RNG
(seed 12345).  N>1 (torchrun, one rank per GPU): every rank validates its own 1 M-configuration
shard (weak scaling) and the validity masks are all-gathered over NCCL — the only collective on
the path.  One JSON line on rank 0 (contract in the task statement):
  value      whole-job configs/s, inputs resident in HBM, CUDA events on the launching stream
  e2e        same metric through the C ABI with pinned HOST buffers (H2D + D2H inside the timing)
  roofline   FP64 pipe (the kernel is compute bound: 53 algorithmic bytes vs ~5e4 flops per
             configuration); achieved = oracle-counted algorithmic flops / kernel time, peak =
             DFMA peak measured live by hppb200_measure_fp64_peak; the HBM view is reported beside it
  cpu_baseline  the oracle (CPU restatement of the reference algorithm) on a bounded sample, warm-up + best of 3
  edges      the metric's other half: validated StraightPath edges/s — Progressive(1e-3) and Dichotomy(0) on 1 M random
             UR5 edges (one GPU's share of BASELINE configs[4]), Discretized(0.05), and BASELINE configs[3]:
             Dichotomy on 1 M local edges of the 30-joint tree model with convex meshes; each with its own FP64
             roofline (oracle-counted flops of THAT batch / CUDA-event time / live DFMA peak), e2e and cpu_baseline
`--impl reference` times the reference algorithm's CPU implementation (oracle port: the reference
binary cannot be built here) on all host threads.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_CONFIGS = 1 << 20          # configurations per GPU per step (BASELINE configs[1])
N_EDGES = 1 << 17            # edges per GPU for the secondary metric (Discretized)
N_EDGES_PROG = 1 << 20       # Progressive, Dichotomy: 1M edges per GPU (BASELINE configs[4]: 10M edges over 8 GPUs)
SEED = 12345
MODEL = "ur5_boxes8.json"
MODEL_HUMANOID = "humanoid_obst16.json"   # BASELINE configs[3]: 30-joint tree, 133 body pairs, convex meshes
HUMANOID_STEP = 0.2                        # local ("RRT-like") edges: |q2 - q1| = 0.2
# Algorithmic flop model (FMA = 2): FK = 126*njoints + 63*ngeoms + 40*njoints (sincos);
# closed-form pair ~110; GJK call = 70 (relative placement + sqrt) + 110 per iteration.
FLOPS_FK = 126 * 6 + 63 * 8 + 40 * 6


def flops_fk(desc):
    return 126 * desc.njoints + 63 * desc.ngeoms + 40 * desc.njoints


WORKLOAD = ("UR5 + 8 boxes: 1M uniform random configurations through CollisionValidation "
            "(FK + narrow phase) per GPU per step (BASELINE configs[1])")
FLOPS_CLOSED = 110
FLOPS_GJK_CALL = 70
FLOPS_GJK_ITER = 110
BYTES_PER_CONFIG = 48 + 1 + 4  # q in, flag + first-pair index out


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-edges", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--full-size", action="store_true",
                    help="also run BASELINE configs[4] at its stated per-GPU size: 12.5 M configurations and 1.25 M edges per rank, one call each")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows = []
        self.proc = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [x.strip() for x in line.split(",")]))

    def stop(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        rows = [r for ts, r in self.rows if t0 <= ts <= t1 + 0.2] or [r for _, r in self.rows]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm = sorted(float(r[0]) for r in rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for k, n in enumerate(names) if any(r[3 + k].lower().startswith("active") for r in rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(rows[0][1]), "reasons": reasons,
                "power_w_max": max(float(r[2]) for r in rows), "samples": len(rows)}


def oracle_counts(orc, q):
    """Algorithmic work of the REFERENCE algorithm (sequential pair loop with early exit) on q."""
    c0 = orc.counters()
    orc.validate_configs(q)
    c1 = orc.counters()
    n = q.shape[0]
    gj = c1["gjk_calls"] - c0["gjk_calls"]
    it = c1["gjk_iterations"] - c0["gjk_iterations"]
    cl = c1["closed_form_calls"] - c0["closed_form_calls"]
    flops = n * FLOPS_FK + cl * FLOPS_CLOSED + gj * FLOPS_GJK_CALL + it * FLOPS_GJK_ITER
    return dict(flops_per_config=flops / n, gjk_calls_per_config=gj / n, gjk_iters_per_config=it / n,
                closed_per_config=cl / n)


def best_of(fn, reps=3):
    """warm-up call, then the best of `reps` timed calls"""
    fn()
    best = 1e30
    for _ in range(reps):
        t = time.perf_counter()
        fn()
        best = min(best, time.perf_counter() - t)
    return best


def cpu_rate_configs(orc, desc, mio, threads, budget_s=4.0):
    """oracle configs/s on a bounded sample of the same workload (first configurations of rank 0's batch)."""
    q = mio.uniform_configs(desc, 20000, seed=SEED)
    t = time.perf_counter()
    orc.validate_configs(q, nthreads=threads)
    rate = 20000 / (time.perf_counter() - t)
    n = int(min(N_CONFIGS, max(20000, rate * budget_s)))
    q = mio.uniform_configs(desc, n, seed=SEED)
    dt = best_of(lambda: orc.validate_configs(q, nthreads=threads))
    return n / dt, n, dt


def cpu_rate_edges(orc, meth, par, q1, q2, L, threads, budget_s=3.0):
    n0 = min(2000, len(L))
    t = time.perf_counter()
    orc.validate_edges(meth, par, q1[:n0], q2[:n0], L[:n0], nthreads=threads)
    rate = n0 / (time.perf_counter() - t)
    n = int(min(len(L), max(n0, rate * budget_s)))
    dt = best_of(lambda: orc.validate_edges(meth, par, q1[:n], q2[:n], L[:n], nthreads=threads))
    return n / dt, n, dt


def oracle_edge_work(orc, desc, meth, par, q1, q2, L, n=20000):
    """Algorithmic flops per edge of the REFERENCE algorithm on the first n edges of the batch (oracle counters)."""
    n = min(n, len(L))
    c0 = orc.counters()
    orc.validate_edges(meth, par, q1[:n], q2[:n], L[:n])
    c1 = orc.counters()
    d = {k: (c1[k] - c0[k]) / n for k in c1}
    flops = (d["samples"] * (flops_fk(desc) + 3 * desc.nq) + d["closed_form_calls"] * FLOPS_CLOSED
             + d["gjk_calls"] * FLOPS_GJK_CALL + d["gjk_iterations"] * FLOPS_GJK_ITER)
    return dict(flops_per_edge=flops, samples_per_edge=d["samples"], gjk_calls_per_edge=d["gjk_calls"],
                gjk_iters_per_edge=d["gjk_iterations"], closed_per_edge=d["closed_form_calls"],
                body_pair_evals_per_edge=d["pair_evals"], counted_on=n)


def run_reference(args):
    """The reference algorithm's own CPU path (oracle port), all host threads, rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import __graft_entry__ as ge
    pkg = ge.load()          # model_io only; no CUDA call is made on this arm
    from oracle import binding as ob
    desc = pkg.ModelDescription.load(MODEL)
    orc = ob.Oracle(desc)
    threads = os.cpu_count() or 1
    per_step = max(1.0, min(10.0, 150.0 / max(1, args.steps + args.warmup)))
    q = pkg.uniform_configs(desc, 20000, seed=SEED)
    t = time.perf_counter()
    orc.validate_configs(q, nthreads=threads)
    rate = 20000 / (time.perf_counter() - t)
    n = int(min(N_CONFIGS, max(20000, rate * per_step)))
    q = pkg.uniform_configs(desc, n, seed=SEED)
    for _ in range(args.warmup):
        orc.validate_configs(q, nthreads=threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        orc.validate_configs(q, nthreads=threads)
    dt = time.perf_counter() - t0
    value = n * args.steps / dt
    sample = "first %d of the %d configurations of rank 0's step batch (seed %d), every step" % (n, N_CONFIGS, SEED)
    line = {
        "impl": "reference", "metric": "validated configs/sec", "value": value, "unit": "configs/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "model": MODEL, "configs_per_gpu_per_step": N_CONFIGS, "seed": SEED,
                   "reference_sample_per_step": n},
        "cpu_baseline": {"value": value, "unit": "configs/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "configs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def run_b200(args):
    import torch
    import torch.distributed as dist
    import __graft_entry__ as ge
    pkg = ge.load()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the engine has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # NCCL's own log (rank counts, NVLS/ring choice) goes to stderr so that stdout stays the one JSON line
        # (the image presets NCCL_DEBUG=VERSION: override it, HPPB200_NCCL_DEBUG chooses another level)
        os.environ["NCCL_DEBUG"] = os.environ.get("HPPB200_NCCL_DEBUG", "INFO")
        os.environ.setdefault("NCCL_DEBUG_SUBSYS", "INIT")
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=dev)

    desc = pkg.ModelDescription.load(MODEL)
    model = pkg.Model(desc)
    ws = pkg.Workspace(model, local_rank)
    stream = torch.cuda.current_stream()
    ws.set_stream(stream.cuda_stream)
    FLAGS_DEV = pkg.DEVICE_POINTERS
    n, nq = N_CONFIGS, model.nq

    # this rank's shard: configurations [rank*n, (rank+1)*n) of the global synthetic stream
    q_host = torch.from_numpy(pkg.uniform_configs(desc, n, seed=SEED, offset=rank * n)).pin_memory()
    q_dev = q_host.to(dev, non_blocking=False)
    valid_d = torch.empty(n, dtype=torch.uint8, device=dev)
    first_d = torch.empty(n, dtype=torch.int32, device=dev)
    reason_d = torch.empty(n, dtype=torch.uint8, device=dev)
    valid_h = torch.empty(n, dtype=torch.uint8).pin_memory()
    first_h = torch.empty(n, dtype=torch.int32).pin_memory()
    reason_h = torch.empty(n, dtype=torch.uint8).pin_memory()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    from hpp_core_b200 import sharding as sh
    gathered = {}
    bits_d = torch.empty((n + 31) // 32, dtype=torch.int32, device=dev)
    allbits_d = torch.empty(world * ((n + 31) // 32), dtype=torch.int32, device=dev)

    # N > 1: the mask gather as ONE kernel over NVLink peer memory (hppb200_exchange_mask: pack + stores into every rank's
    # buffer + epoch flags); checked against the NCCL form once, which stays the fallback (HPPB200_BENCH_PEER=0 forces it)
    px = None
    if world > 1 and os.environ.get("HPPB200_BENCH_PEER", "1") != "0":
        try:
            px = sh.PeerMaskExchange.create(local_rank, n)
        except Exception as exc:  # (every rank takes the same branch only if create() itself agreed: it does, see sharding.py)
            sys.stderr.write("[bench] peer exchange unavailable: %s\n" % exc)
            px = None
        if px is not None:
            ws.validate_configs_ptr(q_dev.data_ptr(), n, valid_d.data_ptr(), first_d.data_ptr(), reason_d.data_ptr(), FLAGS_DEV)
            ref_bits = sh.all_gather_mask_packed(valid_d, unpack=False, bits=bits_d, allbits=allbits_d)[1]
            peer_bits = px.gather(valid_d)
            same = torch.tensor([1 if (not px.timed_out()) and torch.equal(peer_bits, ref_bits) else 0], dtype=torch.int32, device=dev)
            dist.all_reduce(same, op=dist.ReduceOp.MIN)
            if int(same.item()) == 0:
                sys.stderr.write("[bench] peer exchange differs from the NCCL all-gather: NCCL form kept\n")
                px.close()
                px = None

    def step_device():
        ws.validate_configs_ptr(q_dev.data_ptr(), n, valid_d.data_ptr(), first_d.data_ptr(), reason_d.data_ptr(), FLAGS_DEV)
        if world > 1:  # the path's only exchange: gather of the validity masks, bit-packed (1 bit per configuration)
            if px is not None:
                gathered["cfg_bits"] = px.gather(valid_d)
            else:
                gathered["cfg_bits"] = sh.all_gather_mask_packed(valid_d, unpack=False, bits=bits_d, allbits=allbits_d)[1]

    def step_e2e():
        ws.validate_configs_ptr(q_host.data_ptr(), n, valid_h.data_ptr(), first_h.data_ptr(), reason_h.data_ptr(), 0)

    def timed(fn, steps, warmup, all_ranks=True):
        """CUDA-event time of `steps` calls; all_ranks: barrier on both sides and MAX over ranks (every rank must
        call it); all_ranks=False is for the rank-0-only secondary metrics."""
        multi = world > 1 and all_ranks
        for _ in range(warmup):
            fn()
        if multi:
            dist.barrier()
        torch.cuda.synchronize()
        evs = []
        w0 = time.time()
        for _ in range(steps):
            flush.zero_()  # evict inputs/outputs from L2 between timed iterations (not timed)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            fn()
            e1.record(stream)
            evs.append((e0, e1))
        torch.cuda.synchronize()
        if multi:
            dist.barrier()
        w1 = time.time()
        ms = torch.tensor([sum(a.elapsed_time(b) for a, b in evs)], dtype=torch.float64, device=dev)
        if multi:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()), w0, w1

    st0 = ws.stats()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    total_ms, w0, w1 = timed(step_device, args.steps, args.warmup)
    clocks = sampler.stop(w0, w1) if sampler else None
    st1 = ws.stats()
    kernel_ms = st1["last_kernel_ms"]  # device time of the last validate call (events inside the C ABI)
    launches = st1["kernel_launches"] - st0["kernel_launches"] - args.warmup  # kernels inside the timed region
    e2e_ms, _, _ = timed(step_e2e, args.steps, max(3, args.warmup))
    value = world * n * args.steps / (total_ms * 1e-3)
    e2e_value = world * n * args.steps / (e2e_ms * 1e-3)
    # identical results on both paths
    assert torch.equal(valid_h, valid_d.cpu()) and torch.equal(first_h, first_d.cpu())
    if world > 1:  # the gathered (packed) mask holds this rank's flags at its place
        assert torch.equal(sh.unpack_mask(gathered["cfg_bits"], n, world)[rank * n:(rank + 1) * n], valid_d)
        assert px is None or not px.timed_out(), "a peer exchange inside the timed region timed out" 

    # ---- N > 1: every rank validates its own shard of edges (Progressive), validity masks all-gathered ----
    edges_sharded = None
    if not args.no_edges:
        ne = N_EDGES_PROG
        cfg = pkg.uniform_configs(desc, 2 * ne, seed=SEED + 1, offset=rank * 2 * ne)
        s1d = torch.from_numpy(np.ascontiguousarray(cfg[0::2])).to(dev)
        s2d = torch.from_numpy(np.ascontiguousarray(cfg[1::2])).to(dev)
        sLd = torch.from_numpy(ws.straight_lengths(cfg[0::2], cfg[1::2])).to(dev)
        sv = torch.empty(ne, dtype=torch.uint8, device=dev)
        stf = torch.empty(ne, dtype=torch.float64, device=dev)
        stl = torch.empty(ne, dtype=torch.float64, device=dev)
        spr = torch.empty(ne, dtype=torch.int32, device=dev)
        srs = torch.empty(ne, dtype=torch.uint8, device=dev)

        def shard_call():
            ws.validate_edges_ptr(pkg.METHOD_PROGRESSIVE, 1e-3, False, s1d.data_ptr(), s2d.data_ptr(), sLd.data_ptr(), ne,
                                  sv.data_ptr(), stf.data_ptr(), stl.data_ptr(), spr.data_ptr(), srs.data_ptr(), FLAGS_DEV)
            if world > 1:
                gathered["sv_bits"] = sh.all_gather_mask_packed(sv, unpack=False)[1]
        ms_s, _, _ = timed(shard_call, 3, 1)
        sv_all = sh.unpack_mask(gathered["sv_bits"], ne, world) if world > 1 else None
        edges_sharded = {"value": world * ne * 3 / (ms_s * 1e-3), "unit": "edges/s", "edges_per_gpu": ne, "param": 1e-3,
                         "method": "progressive", "ms_per_call": ms_s / 3,
                         "collective": "all_gather of bit-packed validity masks (NCCL, 1 bit per edge)" if world > 1 else "none",
                         "valid_frac_all_ranks": float((sv_all if world > 1 else sv).float().mean().item())}
        del s1d, s2d, sLd, sv, stf, stl, spr, srs, sv_all

    # ---- every rank: node-sharded k-NN (1 M roadmap nodes split by node index, queries replicated), candidates
    #      all-gathered and merged ON THE DEVICE with the reference tie rule (hppb200_knn_merge_shards) ----
    knn_sharded = None
    if not args.no_edges:
        n_nodes, nqr, kk = 1 << 20, 1024, 10
        nb, ne_ = sh.node_owner_ranges(n_nodes, world)[rank]
        nodes_local = pkg.uniform_configs(desc, ne_ - nb, seed=SEED + 2, offset=nb)
        nn_local = pkg.NearestNeighbor(model, local_rank)
        nn_local.append(nodes_local)
        qs_d = torch.from_numpy(pkg.uniform_configs(desc, nqr, seed=SEED + 3)).to(dev)
        li = torch.empty((nqr, kk), dtype=torch.int64, device=dev)
        ld = torch.empty((nqr, kk), dtype=torch.float64, device=dev)
        gi = torch.empty((world, nqr, kk), dtype=torch.int64, device=dev)
        gd = torch.empty((world, nqr, kk), dtype=torch.float64, device=dev)

        def knn_call():
            nn_local.query_ptr(qs_d.data_ptr(), nqr, kk, li.data_ptr(), ld.data_ptr())
            lg = torch.where(li >= 0, li + nb, li)  # global node indices
            if world > 1:
                dist.all_gather_into_tensor(gi.view(-1), lg.view(-1))
                dist.all_gather_into_tensor(gd.view(-1), ld.view(-1))
            else:
                gi[0], gd[0] = lg, ld
            gathered["knn"] = sh.merge_topk_device(gi, gd, kk)
        ms_k, _, _ = timed(knn_call, 5, 2)
        knn_sharded = {"value": nqr * 5 / (ms_k * 1e-3), "unit": "queries/s", "nodes": n_nodes, "nodes_per_gpu": ne_ - nb, "queries": nqr,
                       "k": kk, "ms_per_call": ms_k / 5,
                       "collective": "all_gather of per-rank top-k (distance f64, global index i64) + device merge" if world > 1 else "none"}
        del nn_local

    full = None
    if args.full_size:  # BASELINE configs[4]: 100 M configurations / 10 M edges over 8 GPUs = 12.5 M / 1.25 M per rank, one call each
        nf, nef = 12_500_000, 1_250_000
        qf = torch.empty((nf, nq), dtype=torch.float64, device=dev)
        for c0 in range(0, nf, 1 << 21):
            c1 = min(nf, c0 + (1 << 21))
            qf[c0:c1] = torch.from_numpy(pkg.uniform_configs(desc, c1 - c0, seed=SEED + 7, offset=rank * nf + c0)).to(dev)
        vf = torch.empty(nf, dtype=torch.uint8, device=dev)
        ff = torch.empty(nf, dtype=torch.int32, device=dev)
        rf = torch.empty(nf, dtype=torch.uint8, device=dev)

        def full_cfg():
            ws.validate_configs_ptr(qf.data_ptr(), nf, vf.data_ptr(), ff.data_ptr(), rf.data_ptr(), FLAGS_DEV)
            gathered["vf_bits"] = sh.all_gather_mask_packed(vf, unpack=False)[1]
        ms_fc, _, _ = timed(full_cfg, 3, 1)
        vfrac = float(sh.unpack_mask(gathered["vf_bits"], nf, world).float().mean().item())
        del qf, ff, rf
        cfge = pkg.uniform_configs(desc, 2 * nef, seed=SEED + 8, offset=rank * 2 * nef)
        f1 = torch.from_numpy(np.ascontiguousarray(cfge[0::2])).to(dev)
        f2 = torch.from_numpy(np.ascontiguousarray(cfge[1::2])).to(dev)
        fL = torch.from_numpy(ws.straight_lengths(cfge[0::2], cfge[1::2])).to(dev)
        fv = torch.empty(nef, dtype=torch.uint8, device=dev)
        ftf = torch.empty(nef, dtype=torch.float64, device=dev)
        ftl = torch.empty(nef, dtype=torch.float64, device=dev)
        fpr = torch.empty(nef, dtype=torch.int32, device=dev)
        frs = torch.empty(nef, dtype=torch.uint8, device=dev)

        def full_edges():
            ws.validate_edges_ptr(pkg.METHOD_PROGRESSIVE, 1e-3, False, f1.data_ptr(), f2.data_ptr(), fL.data_ptr(), nef, fv.data_ptr(),
                                  ftf.data_ptr(), ftl.data_ptr(), fpr.data_ptr(), frs.data_ptr(), FLAGS_DEV)
            gathered["fv_bits"] = sh.all_gather_mask_packed(fv, unpack=False)[1]
        ms_fe, _, _ = timed(full_edges, 2, 1)
        full = {"configs": {"value": world * nf * 3 / (ms_fc * 1e-3), "unit": "configs/s", "configs_per_gpu_per_call": nf,
                            "total_configs": world * nf, "ms_per_call": ms_fc / 3, "valid_frac_all_ranks": vfrac},
                "edges": {"value": world * nef * 2 / (ms_fe * 1e-3), "unit": "edges/s", "method": "progressive", "param": 1e-3,
                          "edges_per_gpu_per_call": nef, "total_edges": world * nef, "ms_per_call": ms_fe / 2,
                          "valid_frac_all_ranks": float(sh.unpack_mask(gathered["fv_bits"], nef, world).float().mean().item())},
                "collective": "all_gather of bit-packed validity masks (NCCL)" if world > 1 else "none",
                "note": "BASELINE configs[4]: 100 M configurations / 10 M edges over 8 GPUs; every rank validates its share in ONE call"}
        del f1, f2, fL, fv, ftf, ftl, fpr, frs, vf

    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return

    # ---- rank 0: roofline, CPU baseline, secondary edge metric -------------------------------------
    from oracle import binding as ob
    orc = ob.Oracle(desc)
    threads = os.cpu_count() or 1
    sub = q_host.numpy()[:50000]
    work = oracle_counts(orc, sub)
    ov, op, _ = orc.validate_configs(sub)
    parity_ok = bool(np.array_equal(ov, valid_h.numpy()[:50000]) and np.array_equal(op, first_h.numpy()[:50000]))
    peak = pkg.measure_fp64_peak(local_rank)
    per_launch_s = (total_ms * 1e-3) / args.steps if world == 1 else kernel_ms * 1e-3
    achieved = work["flops_per_config"] * n / per_launch_s / 1e12
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    traffic = {}
    try:  # dram__bytes_read.sum + dram__bytes_write.sum per launch from an `ncu --set full` capture of THIS build
        traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
    except Exception:
        pass
    roofline = {"bound": "fp64", "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
                # dram__bytes_read.sum + dram__bytes_write.sum per launch, ncu --set full, 1 M configurations
                # (profiles/r1b_configPass_gjkJobs_ncu_full.txt): the placement scratch written once per configuration;
                # removing it was measured slower
                "traffic": traffic.get("configPassKernel_dram_bytes") if n == (1 << 20) else None,
                "traffic_unit": "B per launch of configPassKernel", "traffic_source": traffic.get("source"),
                "peak_source": "hppb200_measure_fp64_peak (DFMA chains, measured live)",
                "kernel": "configPassKernel + gjkJobsKernel", "algorithmic_flops_per_config": work["flops_per_config"],
                "hbm": {"achieved": BYTES_PER_CONFIG * n / per_launch_s / 1e9, "peak": hbm_peak, "unit": "GB/s",
                        "frac": BYTES_PER_CONFIG * n / per_launch_s / 1e9 / hbm_peak,
                        "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback"},
                "device_pair_tests_per_config": (st1["pair_tests"] - st0["pair_tests"]) / (n * (args.steps + args.warmup)),
                "reference_pair_tests_per_config": work["gjk_calls_per_config"] + work["closed_per_config"]}
    cpu = None
    if not args.no_cpu:
        rate, ns, dt = cpu_rate_configs(orc, desc, pkg, threads)
        cpu = {"value": rate, "unit": "configs/s", "cores": threads, "kind": "port",
               "sample": "first %d of the %d configurations of the step batch, %.1f s" % (ns, n, dt)}
    edges = None
    if not args.no_edges:
        def edge_suite(label, e_desc, e_model, e_ws, e_orc, meth, par, q1n, q2n, ne):
            """device (inputs in HBM), e2e (host buffers, all five outputs copied back), roofline, cpu_baseline"""
            e1h = torch.from_numpy(np.ascontiguousarray(q1n)).pin_memory()
            e2h = torch.from_numpy(np.ascontiguousarray(q2n)).pin_memory()
            L_h = torch.from_numpy(e_ws.straight_lengths(e1h.numpy(), e2h.numpy())).pin_memory()
            e1d, e2d, Ld = e1h.to(dev), e2h.to(dev), L_h.to(dev)
            ev_d = torch.empty(ne, dtype=torch.uint8, device=dev)
            tf_d = torch.empty(ne, dtype=torch.float64, device=dev)
            tl_d = torch.empty(ne, dtype=torch.float64, device=dev)
            pr_d = torch.empty(ne, dtype=torch.int32, device=dev)
            rs_d = torch.empty(ne, dtype=torch.uint8, device=dev)
            ev_h = torch.empty(ne, dtype=torch.uint8).pin_memory()
            tf_h = torch.empty(ne, dtype=torch.float64).pin_memory()
            tl_h = torch.empty(ne, dtype=torch.float64).pin_memory()
            pr_h = torch.empty(ne, dtype=torch.int32).pin_memory()
            rs_h = torch.empty(ne, dtype=torch.uint8).pin_memory()

            def dev_call():
                e_ws.validate_edges_ptr(meth, par, False, e1d.data_ptr(), e2d.data_ptr(), Ld.data_ptr(), ne, ev_d.data_ptr(),
                                        tf_d.data_ptr(), tl_d.data_ptr(), pr_d.data_ptr(), rs_d.data_ptr(), FLAGS_DEV)

            def host_call():
                e_ws.validate_edges_ptr(meth, par, False, e1h.data_ptr(), e2h.data_ptr(), L_h.data_ptr(), ne, ev_h.data_ptr(),
                                        tf_h.data_ptr(), tl_h.data_ptr(), pr_h.data_ptr(), rs_h.data_ptr(), 0)
            s0 = e_ws.stats()
            ms_d, _, _ = timed(dev_call, 3, 1, all_ranks=False)
            s1 = e_ws.stats()
            ms_h, _, _ = timed(host_call, 3, 1, all_ranks=False)
            same = bool(torch.equal(ev_h, ev_d.cpu()) and torch.equal(pr_h, pr_d.cpu()))
            out = {"value": ne * 3 / (ms_d * 1e-3), "unit": "edges/s", "n_edges": ne, "param": par, "model": e_desc.name,
                   "ms_per_call": ms_d / 3,
                   "e2e": {"value": ne * 3 / (ms_h * 1e-3), "unit": "edges/s", "ms_per_call": ms_h / 3,
                           "h2d_bytes_per_step": ne * (2 * e_model.nq + 1) * 8, "d2h_bytes_per_step": ne * (1 + 8 + 8 + 4 + 1)},
                   "device_and_host_paths_identical": same,
                   "valid_frac": float(ev_h.numpy()[:ne].mean()),
                   "kernel_launches_per_call": (s1["kernel_launches"] - s0["kernel_launches"]) / 4,
                   "samples_per_edge": (s1["edge_samples"] - s0["edge_samples"]) / (4 * ne),
                   "pair_tests_per_edge": (s1["pair_tests"] - s0["pair_tests"]) / (4 * ne)}
            if not args.no_cpu:
                q1s, q2s, Ls = e1h.numpy(), e2h.numpy(), L_h.numpy()
                work = oracle_edge_work(e_orc, e_desc, meth, par, q1s, q2s, Ls)
                ach = work["flops_per_edge"] * ne / (ms_d * 1e-3 / 3) / 1e12
                out["roofline"] = {"bound": "fp64", "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak,
                                   "algorithmic_flops_per_edge": work["flops_per_edge"], "work": work,
                                   "peak_source": "hppb200_measure_fp64_peak (DFMA chains, measured live)",
                                   "kernel_shares": traffic.get(label + "_kernel_shares"),
                                   "hbm": {"achieved": ne * ((2 * e_model.nq + 1) * 8 + 22) / (ms_d * 1e-3 / 3) / 1e9,
                                           "peak": hbm_peak, "unit": "GB/s"}}
                # parity of the leading edges against the oracle (the same call the counters came from)
                chk = min(ne, 20000)
                oe = e_orc.validate_edges(meth, par, q1s[:chk], q2s[:chk], Ls[:chk])
                out["parity_vs_oracle"] = bool(np.array_equal(oe["valid"], ev_h.numpy()[:chk])
                                               and np.array_equal(oe["t_fail"], tf_h.numpy()[:chk], equal_nan=True)
                                               and np.array_equal(oe["t_last_valid"], tl_h.numpy()[:chk])
                                               and np.array_equal(oe["pair"], pr_h.numpy()[:chk]))
                rate, ns, dt = cpu_rate_edges(e_orc, meth, par, q1s, q2s, Ls, threads)
                out["cpu_baseline"] = {"value": rate, "unit": "edges/s", "cores": threads, "kind": "port",
                                       "sample": "first %d of the %d edges, warm-up + best of 3 (%.1f s)" % (ns, ne, dt)}
            return out

        cfg = pkg.uniform_configs(desc, 2 * N_EDGES_PROG, seed=SEED + 1, offset=0)
        res = {}
        for name, meth, par in (("progressive", pkg.METHOD_PROGRESSIVE, 1e-3), ("discretized", pkg.METHOD_DISCRETIZED, 0.05),
                                ("dichotomy", pkg.METHOD_DICHOTOMY, 0.0)):
            ne = N_EDGES if name == "discretized" else N_EDGES_PROG  # leading `ne` edges of the same stream
            res[name] = edge_suite(name, desc, model, ws, orc, meth, par, cfg[0:2 * ne:2], cfg[1:2 * ne:2], ne)
        # BASELINE configs[3]: 30-joint tree with convex meshes, Dichotomy, 1 M local edges
        hdesc = pkg.ModelDescription.load(MODEL_HUMANOID)
        hmodel = pkg.Model(hdesc)
        hws = pkg.Workspace(hmodel, local_rank)
        hws.set_stream(stream.cuda_stream)
        horc = ob.Oracle(hdesc)
        hq1 = pkg.uniform_configs(hdesc, N_EDGES_PROG, seed=SEED + 5)
        hd = pkg.uniform_configs(hdesc, N_EDGES_PROG, seed=SEED + 6) - hq1
        hq2 = pkg.normalize_quaternions(hdesc, np.clip(hq1 + HUMANOID_STEP * hd / np.linalg.norm(hd, axis=1, keepdims=True),
                                                       hdesc.q_lower, hdesc.q_upper))
        for name, meth, par in (("dichotomy_humanoid", pkg.METHOD_DICHOTOMY, 0.0), ("progressive_humanoid", pkg.METHOD_PROGRESSIVE, 1e-3)):
            res[name] = edge_suite(name, hdesc, hmodel, hws, horc, meth, par, hq1, hq2, N_EDGES_PROG)
            res[name]["edges"] = "local: |q2 - q1| = %g in configuration space (SURVEY 8d C4)" % HUMANOID_STEP
        del hws, hmodel
        edges = res
    knn = None
    if not args.no_edges:
        # secondary metric: nearestNeighbor::Basic over a 1M-node roadmap mirror (BASELINE configs[2]); wall clock
        # through the C ABI with host pointers (query upload + result download inside), median of 20 calls
        nn = pkg.NearestNeighbor(model, local_rank)
        os.environ["HPPB200_KNN_TC"] = "0"  # a second mirror restricted to the exact FP64 scan, for comparison
        nn_exact = pkg.NearestNeighbor(model, local_rank)
        del os.environ["HPPB200_KNN_TC"]
        n_nodes = 1 << 20
        nn.append(pkg.uniform_configs(desc, n_nodes, seed=SEED + 2))
        nn_exact.append(pkg.uniform_configs(desc, n_nodes, seed=SEED + 2))
        knn = {"nodes": n_nodes, "unit": "queries/s", "bytes_per_node": 8 * nq}
        tmem_peak = pkg.measure_tmem_read_peak(local_rank)  # TB/s (the library reports TB/s)
        if "bf16_tflops" in peaks:
            tf32_peak = (peaks["bf16_tflops"] / 2.0, "MEASURED_PEAKS.json bf16_tflops (burst) / 2: kind::tf32 runs at half the bf16 rate")
        else:
            tf32_peak = (2250.0 / 2.0, "B200_PROFILING.md fallback: nominal 2.25 PFLOP/s bf16 / 2")
        for label, obj, nqr, kk, reps in (("q1_k1", nn, 1, 1, 20), ("q1_k10", nn, 1, 10, 20), ("q1024_k1", nn, 1024, 1, 9),
                                          ("q1024_k10", nn, 1024, 10, 9), ("q8192_k10", nn, 8192, 10, 5),
                                          ("q1024_k1_exact_scan", nn_exact, 1024, 1, 5),
                                          ("q1024_k10_exact_scan", nn_exact, 1024, 10, 3)):
            qs = pkg.uniform_configs(desc, nqr, seed=SEED + 3)
            ref_idx, ref_dd = obj.query(qs, kk)
            ts = []
            for _ in range(reps):
                t0 = time.perf_counter()
                idx, dd = obj.query(qs, kk)
                ts.append(time.perf_counter() - t0)
            tm = sorted(ts)[len(ts) // 2]
            st = obj.tc_stats()
            knn[label] = {"value": nqr / tm, "ms_per_call": tm * 1e3}
            if st["used"]:  # tcgen05 TF32 candidate pass + exact FP64 re-rank (kernels_knn_tc.cu)
                knn[label].update({"path": "tcgen05 tf32 candidate pass + fp64 re-rank", "prefix_stages": st["stages"],
                                   "candidate_groups_of_8_per_query": st["candidates"] / nqr,
                                   "queries_redone_by_exact_scan": st["redone"]})
                # the whole-roadmap filter pass (knnTcFilterKernel), CUDA events inside the library.  Two ceilings, both
                # MEASURED: the tensor pipe (2 x 8 kc flop per (query, node) pair in TF32 = half the bf16 rate; bf16 burst
                # from MEASURED_PEAKS.json) and the TMEM read path (one FP32 accumulator word per pair leaves tensor memory
                # through tcgen05.ld; hppb200_measure_tmem_read_peak, the filter's own access shape).  The binding one is
                # the roofline; the other is reported beside it.
                if st["filter_ms"] > 0:
                    sec = st["filter_ms"] * 1e-3
                    kc = (3 * nq + 6 + 7) // 8
                    tf = 2.0 * 8 * kc * st["filter_pairs"] / sec / 1e12
                    tb = 4.0 * st["filter_pairs"] / sec / 1e12
                    tf_peak, tf_src = tf32_peak
                    r_tensor = {"bound": "tensor", "achieved": tf, "peak": tf_peak, "unit": "TFLOP/s", "frac": tf / tf_peak,
                                "peak_source": tf_src, "flop_per_pair": 16 * kc}
                    r_tmem = {"bound": "tmem-read", "achieved": tb, "peak": tmem_peak, "unit": "TB/s", "frac": tb / tmem_peak,
                              "peak_source": "measured: hppb200_measure_tmem_read_peak (16 warps x tcgen05.ld.32x32b.x32, CUDA events)",
                              "bytes_per_pair": 4}
                    first, second = (r_tensor, r_tmem) if r_tensor["frac"] >= r_tmem["frac"] else (r_tmem, r_tensor)
                    knn[label]["filter_kernel"] = {"ms": st["filter_ms"], "pairs_per_s": st["filter_pairs"] / sec,
                                                   "roofline": first, "other_ceiling": second}
            else:
                knn[label].update({"path": "exact fp64 scan", "scan_passes": 1 if kk <= 32 else kk,
                                   "node_stream_gbs": n_nodes * 8 * nq * (1 if kk <= 32 else kk) * ((nqr + 3) // 4) / tm / 1e9})
            if label.endswith("_exact_scan"):  # both paths return the same bits
                ti, td = nn.query(qs, kk)
                knn[label]["identical_to_tensor_core_path"] = bool(np.array_equal(ti, idx) and np.array_equal(td, dd))
        del nn_exact
        # the single-query kernel alone (BiRRT's nearestNode): CUDA events on the mirror's stream around the one launch,
        # against the HBM copy peak: the node stream (8 nq bytes per node) is all it reads
        os.environ["HPPB200_KNN_TIMING"] = "1"
        nn_t = pkg.NearestNeighbor(model, local_rank)
        del os.environ["HPPB200_KNN_TIMING"]
        nn_t.append(pkg.uniform_configs(desc, n_nodes, seed=SEED + 2))
        q1 = pkg.uniform_configs(desc, 1, seed=SEED + 3)
        for kk in (1, 10):
            ms = []
            for _ in range(30):
                nn_t.query(q1, kk)
                t_ms, nbytes = nn_t.latency_time()
                if t_ms > 0:
                    ms.append(t_ms)
            if ms:
                med = sorted(ms)[len(ms) // 2]
                gbs = nbytes / (med * 1e-3) / 1e9
                knn["q1_k%d" % kk]["kernel"] = {"ms": med, "roofline": {"bound": "hbm", "achieved": gbs, "peak": hbm_peak, "unit": "GB/s",
                                                                          "frac": gbs / hbm_peak, "bytes_per_node": 8 * nq,
                                                                          "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback"}}
        del nn_t
        if not args.no_cpu:
            orc.knn_append(pkg.uniform_configs(desc, n_nodes, seed=SEED + 2))
            qs = pkg.uniform_configs(desc, 64, seed=SEED + 3)
            t0 = time.perf_counter()
            oi, od = orc.knn_query(qs, 10, nthreads=threads)
            dt = time.perf_counter() - t0
            gi, gd = nn.query(qs, 10)
            knn["cpu_baseline"] = {"value": 64 / dt, "unit": "queries/s", "cores": threads, "kind": "port",
                                   "sample": "64 queries, k = 10, %.2f s" % dt}
            knn["parity_64_queries_k10"] = bool(np.array_equal(oi, gi) and np.array_equal(od, gd))

    # ---- BASELINE configs[2]: BiRRT over a 1 M-node roadmap: planner steps/s through the unchanged one-item-per-call
    #      interfaces (the honest number: every step is 2 nearest-node queries, up to 3 path validations, 1 M-node
    #      scans) and through the batched step (one k-NN batch + one edge batch per phase), oracle backend beside it ----
    birrt = None
    if not args.no_edges:
        from hpp_core_b200 import planners as pl
        n_pre = 1 << 19  # nodes per tree: 2 x 512 k = 1 M roadmap nodes

        def prefilled(backend):
            q0, q1 = pkg.uniform_configs(desc, 1, seed=SEED + 10)[0], pkg.uniform_configs(desc, 1, seed=SEED + 11)[0]
            planner = pl.BiRRTPlanner(backend, desc, q0, q1, seed=SEED + 12)
            rmap = planner.roadmap
            rmap.extend(pkg.uniform_configs(desc, n_pre - 1, seed=SEED + 13), np.zeros(n_pre - 1, np.int64))       # start tree, label 0
            rmap.extend(pkg.uniform_configs(desc, n_pre - 1, seed=SEED + 14), np.ones(n_pre - 1, np.int64))        # goal tree, label 1
            return planner

        def step_rate(planner, nsteps, batch):
            ts = []
            for _ in range(nsteps):
                t0 = time.perf_counter()
                if batch:
                    planner.one_step_batch(batch)
                else:
                    planner.one_step()
                ts.append(time.perf_counter() - t0)
            med = sorted(ts)[len(ts) // 2]  # (median: the first component merge relabels 512 k host labels once)
            return (batch or 1) / med, med

        eng = prefilled(pl.EngineBackend(pkg, model, local_rank, pkg.METHOD_PROGRESSIVE, 1e-3))
        step_rate(eng, 5, 0)
        r1, t1 = step_rate(eng, 60, 0)
        step_rate(eng, 1, 256)
        rb, tb = step_rate(eng, 7, 256)
        birrt = {"roadmap_nodes": eng.roadmap.n, "unit": "planner steps/s",
                 "single_item_calls": {"value": r1, "ms_per_step": t1 * 1e3,
                                       "note": "BiRRTPlanner::oneStep as the reference drives it: one configuration / one edge per call"},
                 "batched_step_256": {"value": rb, "ms_per_batch": tb * 1e3, "extensions_per_batch": 256}}
        del eng
        if not args.no_cpu:
            sys.path.insert(0, os.path.join(ROOT, "tests"))
            from oracle_backend import OracleBackend
            cpu_pl = prefilled(OracleBackend(ob.Oracle(desc), pkg.METHOD_PROGRESSIVE, 1e-3))
            step_rate(cpu_pl, 2, 0)
            c1, ct1 = step_rate(cpu_pl, 12, 0)
            cb, ctb = step_rate(cpu_pl, 2, 256)
            birrt["cpu_baseline"] = {"single_item_calls": c1, "batched_step_256": cb, "unit": "planner steps/s", "cores": threads,
                                     "kind": "port", "sample": "12 single steps / 2 batched steps over the same pre-filled roadmap"}
            del cpu_pl

    if knn_sharded is not None and world > 1:  # parity of the sharded result with the unsharded mirror (rank 0)
        whole = pkg.NearestNeighbor(model, local_rank)
        whole.append(pkg.uniform_configs(desc, 1 << 20, seed=SEED + 2))
        wi, wd = whole.query(pkg.uniform_configs(desc, 1024, seed=SEED + 3), 10)
        gi_, gd_ = gathered["knn"]
        knn_sharded["identical_to_unsharded"] = bool(np.array_equal(gi_.cpu().numpy(), wi) and np.array_equal(gd_.cpu().numpy(), wd))
        del whole

    line = {
        "metric": "validated configs/sec", "value": value, "unit": "configs/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "model": MODEL, "configs_per_gpu_per_step": n, "seed": SEED, "pairs": int(len(model.config_pairs()[0])),
                   "l2": "flushed between timed steps (256 MiB memset, untimed)",
                   "collective": ("none" if world == 1 else
                                  "bit-packed validity masks stored into every rank's buffer over NVLink peer memory by one kernel (hppb200_exchange_mask), 1 bit per configuration"
                                  if px is not None else "all_gather of bit-packed validity masks (NCCL, 1 bit per configuration)")},
        "e2e": {"value": e2e_value, "unit": "configs/s", "h2d_bytes_per_step": n * nq * 8, "d2h_bytes_per_step": n * 6,
                "ms_per_step": e2e_ms / args.steps},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": roofline,
        "cpu_baseline": cpu,
        "edges": edges,
        "edges_sharded": edges_sharded,
        "knn_sharded": knn_sharded,
        "full_size": full,
        "birrt": birrt,
        "knn": knn,
        "parity_vs_oracle_50k": parity_ok,
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
