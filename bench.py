#!/usr/bin/env python
"""bench.py -- seconds per tree and histogram rows x features / s of the GBDT hot path.

A step = one boosting iteration of the hot path: the round's (g,h) pairs are pushed to
the tree maker and ITreeMaker.make() builds one tree (histograms, subtraction, split
search, row partition).  Default workload = BASELINE.json configs[1]: Higgs shape,
11 M x 28 dense, 255 bins, depth 8, synthetic.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

N > 1 is launched by torchrun, one rank per GPU; rows are sharded (each rank holds
--rows rows: weak scaling) and the per-node histogram reduce-scatter + split allgather
run over NCCL.  `value` is timed with the round's gradients already in HBM; `e2e`
includes the host->device copy of the gradients from pinned memory and the tree coming
back to the host.  --impl reference times the CPU restatement of the reference's Java
path (oracle/, "port": there is no JVM in this image) on the host cores.
"""
import argparse
import json
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

METRIC = "hist_rows_x_features_per_sec"
UNIT = "rows*features/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--rows", type=int, default=11_000_000, help="rows per GPU")
    ap.add_argument("--features", type=int, default=28)
    ap.add_argument("--bins", type=int, default=255)
    ap.add_argument("--depth", type=int, default=8)
    ap.add_argument("--policy", default="level", choices=["level", "loss"])
    ap.add_argument("--leaves", type=int, default=-1)
    ap.add_argument("--cpu-rows", type=int, default=2_000_000, help="rows of the bounded CPU-baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-loss-wise", action="store_true", help="skip the extra loss-wise 255-leaf measurement")
    ap.add_argument("--no-parity-check", action="store_true", help="skip the self-check after the timed region")
    ap.add_argument("--no-other-configs", action="store_true", help="skip BASELINE configs[2..4] (other_configs)")
    ap.add_argument("--no-strong", action="store_true", help="N > 1: skip the strong-scaling extra (strong_11M)")
    ap.add_argument("--c3-rows", type=int, default=12_500_000, help="rows per GPU of configs[2]")
    ap.add_argument("--c5-rows", type=int, default=20_000_000, help="rows per GPU of configs[4]")
    ap.add_argument("--c4-rows", type=int, default=6_250_000, help="rows per GPU of configs[3]")
    ap.add_argument("--uniform-bins", action="store_true", help="experiment: every feature uniform over its bins")
    return ap.parse_args()


def tree_kwargs(a):
    return dict(max_depth=a.depth, max_leaf_cnt=a.leaves, grow_policy=a.policy, min_child_hessian_sum=100.0,
                min_split_loss=0.0, l1=0.0, l2=0.0, learning_rate=0.1, min_split_samples=-1)


import workloads as wl          # tests/workloads.py: the synthetic inputs of the BASELINE shapes

synth_bins = wl.c2_higgs        # (n, F, nb, seed, uniform=False) -> uint8 slots [n,F], labels
sigmoid_grads = wl.sigmoid_grads


class ClockSampler:
    """nvidia-smi clocks and throttle reasons while the steps run.  The process is started (and NVML
    initialised -- which can stall CUDA calls of other processes for tens of ms) well before the
    timed region; only samples taken after mark() are used."""
    Q = ("timestamp,index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu):
        self.proc = None
        self.t_mark = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(gpu), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "250"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None

    def mark(self):
        import datetime
        self.t_mark = datetime.datetime.now()

    def stop(self):
        import datetime
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except Exception:
            self.proc.kill()
            out = ""
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in out.strip().splitlines():
            c = [t.strip() for t in line.split(",")]
            if len(c) < 10:
                continue
            try:
                ts = datetime.datetime.strptime(c[0], "%Y/%m/%d %H:%M:%S.%f")
                if self.t_mark is not None and ts < self.t_mark:
                    continue
                sm.append(float(c[2])); mx.append(float(c[3]))
            except ValueError:
                continue
            for nme, v in zip(names, c[6:10]):
                if v.lower().startswith("active"):
                    reasons.add(nme)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                "samples": len(sm)}


def measured_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic():
    """DRAM bytes per histogram-build launch from the committed ncu launch list (tools/summarize_ncu.py) -- reported
    only while the kernels' source is byte for byte what was profiled (the summary records its hash); else null."""
    path = os.path.join(ROOT, "profiles", "hist_build_ncu.json")
    try:
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import summarize_ncu
        d = json.load(open(path))
        if d.get("kernel_source_sha256_16") != summarize_ncu.kernel_source_hash():
            return None, "profiles/hist_build_ncu.json was taken from an older build of the kernel: not reported"
        return float(d["dram_bytes_per_launch"]), "ncu launch list %s (commit %s), same kernel source" % (d.get("source"), d.get("summarised_at_commit"))
    except Exception:
        return None, "no capture"


def cpu_baseline(a, xs, ys, steps=1, threads=None):
    """The CPU restatement of the reference's path (oracle, "port") on the host cores: one
    emulated worker per core, each with its own row shard, as the Java trainer runs one
    mp4j rank per thread.  Bounded sample: the first --cpu-rows rows of the same workload."""
    import oracle_binding as ob
    cores = threads or os.cpu_count() or 1
    n = min(a.cpu_rows, xs.shape[0])
    x = xs[:n].astype(np.int32)
    y = ys[:n]
    start = (np.arange(a.features + 1) * a.bins).astype(np.int32)
    p = ob.make_params(**tree_kwargs(a))
    score = np.zeros(n, np.float32)
    gp = np.empty(2 * n, np.float32)
    cells = secs = 0.0
    nodes = 0
    for _ in range(steps):
        sigmoid_grads(score, y, gp)
        t, pos, info = ob.make_tree(p, x, gp, 0, start, W=cores, threads=cores)
        cells += info["hist_cells"]; secs += info["t_total"]; nodes = t.n_nodes
        score += t.node_info[pos]
    return {"value": cells / secs, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": "%d trees on the first %d of %d rows x %d features, W=%d emulated workers"
                      % (steps, n, xs.shape[0], a.features, cores),
            "seconds_per_tree_on_sample": secs / steps, "nodes": nodes}


def _tree_digest(tree, fields):
    import hashlib
    hd = hashlib.sha256()
    hd.update(np.int64(tree.n_nodes).tobytes())
    for name, _ in fields:
        hd.update(np.ascontiguousarray(getattr(tree, name)).tobytes())
    return np.frombuffer(hd.digest(), dtype=np.uint8).copy()


def parity_check(a, yb, torch, dist, world, rank, local, fresh_nccl_id, last_tree, rows_small=1_000_000):
    """Outside every timer.  (1) every rank holds the same tree (all fields, byte for byte) and its root
    counts the global rows; (2) on a second, small tree maker (rows_small rows per rank, same generator):
    the N-rank tree and the concatenated per-row leaves equal the tree ONE GPU builds on the concatenated
    rows (a world = 1 handle on rank 0), bit for bit, level-wise depth 8 and loss-wise 255 leaves; and
    that tree equals the CPU oracle's (W = N emulated workers): structure, split features / intervals,
    counts and per-row leaves exactly, values to 1e-5.  Returns a dict; "ok" False makes bench.py exit 1."""
    res = {"ok": True, "checks": [], "rows_per_rank_small": rows_small}

    def note(name, ok, detail=None):
        res["checks"].append({"name": name, "ok": bool(ok), **({"detail": detail} if detail else {})})
        res["ok"] = res["ok"] and bool(ok)

    def all_ok(flag):                       # every rank learns whether any rank failed
        if world == 1:
            return flag
        t = torch.tensor([1 if flag else 0], dtype=torch.int32, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        return bool(t.item())

    # (1) the timed workload's last tree
    dig = torch.from_numpy(_tree_digest(last_tree, yb.Tree.FIELDS)).cuda()
    if world > 1:
        allg = [torch.empty_like(dig) for _ in range(world)]
        dist.all_gather(allg, dig)
        same = all(bool(torch.equal(g, allg[0])) for g in allg)
    else:
        same = True
    if rank == 0:
        note("all_ranks_hold_the_same_tree", same)
        note("root_counts_global_rows", int(last_tree.sample_cnt[0]) == a.rows * world,
             "%d vs %d" % (int(last_tree.sample_cnt[0]), a.rows * world))

    # (2) small second tree maker
    n, F = rows_small, a.features
    start = (np.arange(F + 1) * a.bins).astype(np.int32)
    xs, ys = synth_bins(n, F, a.bins, 4321 + rank)
    gp = np.empty(2 * n, np.float32)
    sigmoid_grads(np.zeros(n, np.float32), ys, gp)
    settings = {"level_depth8": dict(tree_kwargs(a), grow_policy="level", max_depth=8, max_leaf_cnt=-1),
                "loss_255": dict(tree_kwargs(a), grow_policy="loss", max_depth=-1, max_leaf_cnt=255)}
    for name, kw in settings.items():
        tm = yb.GpuTreeMaker(yb.make_params(**kw), start, n, device=local, rank=rank, world=world,
                             nccl_id=fresh_nccl_id())
        tm.upload_rows(xs.astype(np.int32))
        tm.set_grads(gp)
        tree = tm.make(0)
        pos = tm.get_positions()
        tm.close()
        if world > 1:
            pos_t = torch.from_numpy(pos).cuda()
            gathered = [torch.empty_like(pos_t) for _ in range(world)] if rank == 0 else None
            dist.gather(pos_t, gathered, dst=0)
        if rank == 0:
            # the same rows on ONE GPU, and through the CPU oracle
            xa = [xs] + [synth_bins(n, F, a.bins, 4321 + r)[0] for r in range(1, world)]
            ya = [ys] + [synth_bins(n, F, a.bins, 4321 + r)[1] for r in range(1, world)]
            x_all = np.concatenate(xa).astype(np.int32)
            gp_all = np.empty(2 * n * world, np.float32)
            sigmoid_grads(np.zeros(n * world, np.float32), np.concatenate(ya), gp_all)
            pos_all = pos if world == 1 else np.concatenate([g.cpu().numpy() for g in gathered])
            if world > 1:
                one = yb.GpuTreeMaker(yb.make_params(**kw), start, n * world, device=local)
                one.upload_rows(x_all)
                one.set_grads(gp_all)
                t1 = one.make(0)
                p1 = one.get_positions()
                one.close()
                eq = t1.n_nodes == tree.n_nodes and all(
                    np.array_equal(getattr(t1, f), getattr(tree, f), equal_nan=True) for f, _ in yb.Tree.FIELDS)
                note(name + ":%d_ranks_equal_one_gpu_bit_for_bit" % world, eq and np.array_equal(p1, pos_all))
            import oracle_binding as ob
            from gpu_common import assert_same_tree
            o, opos, _ = ob.make_tree(ob.make_params(**kw), x_all, gp_all, 0, start, W=max(world, 1),
                                      threads=os.cpu_count() or 1)
            try:
                assert_same_tree(tree, o)
                ok, why = np.array_equal(pos_all, opos), "per-row leaves"
            except AssertionError as e:
                ok, why = False, str(e)[:200]
            note(name + ":equals_cpu_oracle_W%d" % world, ok, None if ok else why)
            res.setdefault("nodes", {})[name] = int(tree.n_nodes)
    res["ok"] = all_ok(res["ok"])
    return res


def _gen_c3(torch, n, F, nb, seed):
    """configs[2] shape on the device: uniform slots, y = sum_j a_j slot_j + N(0,1), l2 gradients at score 0."""
    g = torch.Generator(device="cuda"); g.manual_seed(seed)
    x = torch.randint(0, nb, (n, F), dtype=torch.uint8, device="cuda", generator=g)
    a = torch.randn(F, device="cuda", generator=g) * (torch.rand(F, device="cuda", generator=g) < 0.25) / nb
    y = torch.zeros(n, device="cuda")
    for f in torch.nonzero(a).flatten().tolist():
        y += x[:, f].float() * a[f]
    y += torch.randn(n, device="cuda", generator=g)
    gp = torch.stack([-y, torch.ones_like(y)], 1).reshape(-1)
    return x.cpu().numpy(), gp.cpu().numpy()


def _gen_c5(torch, n, F, nb, K, seed):
    """configs[4] shape on the device: uniform slots, K classes, softmax gradients at score 0."""
    g = torch.Generator(device="cuda"); g.manual_seed(seed)
    x = torch.randint(0, nb, (n, F), dtype=torch.uint8, device="cuda", generator=g)
    w = torch.randn(8, K, device="cuda", generator=g) * 3
    z = (x[:, :8].float() / nb - 0.5) @ w
    z += -torch.log(-torch.log(torch.rand(n, K, device="cuda", generator=g).clamp_min(1e-9)))
    label = z.argmax(1)
    p = 1.0 / K
    gp = torch.empty(n, K, 2, device="cuda")
    gp[:, :, 0] = p
    gp[torch.arange(n, device="cuda"), label, 0] = p - 1.0
    gp[:, :, 1] = 2.0 * p * (1.0 - p)
    return x.cpu().numpy(), gp.reshape(-1).cpu().numpy()


def _gen_c4(torch, n, F, density, seed):
    """configs[3] shape on the device: k = F*density stored entries per row, one per stratum of F/k features,
    slots 1..b_f-1 with b_f in [2,255], default slot 0; sigmoid gradients at score 0 from labels that
    depend on a few features."""
    g = torch.Generator(device="cuda"); g.manual_seed(seed)
    k = max(1, int(round(F * density)))
    width = F // k
    nbs = torch.randint(2, 256, (F,), device="cuda", generator=g)
    fids = torch.arange(k, device="cuda")[None, :] * width + torch.randint(0, width, (n, k), device="cuda", generator=g)
    nbf = nbs[fids]
    slots = 1 + (torch.rand(n, k, device="cuda", generator=g) * (nbf - 1)).long().clamp_max(253)
    slots = torch.minimum(slots, nbf - 1)
    w = torch.randn(F, device="cuda", generator=g) * (torch.rand(F, device="cuda", generator=g) < 0.05)
    z = (w[fids] * (slots.float() / nbf.float())).sum(1)
    y = (torch.rand(n, device="cuda", generator=g) < torch.sigmoid(3 * z)).float()
    gp = torch.stack([0.5 - y, torch.full_like(y, 0.25)], 1).reshape(-1)
    start = np.concatenate([[0], np.cumsum(nbs.cpu().numpy())]).astype(np.int32)
    row_ptr = np.arange(n + 1, dtype=np.int64) * k
    return start, row_ptr, fids.reshape(-1).int().cpu().numpy(), slots.reshape(-1).int().cpu().numpy(), gp.cpu().numpy()


def other_configs(a, yb, torch, dist, world, rank, local, fresh_nccl_id, barrier, peak):
    """BASELINE.json configs[2..4] on this run's GPUs (north_star: "throughput on synthetic data of the named
    shapes at 1, 2, 4 and 8 GPUs"): per-GPU shard of each shape generated on the device, a few trees timed
    like the headline workload (gradients resident in HBM, max over ranks), each with the roofline of its
    own dominant kernel.  Parity for these shapes: tests/test_gpu_fullsize.py (2 M / 1 M / 1 M rows)."""
    out = {}

    def timed(tm, gids, warm, steps):
        cells = rows = rows_idx = built = hist_ms = t_sum = 0.0
        nodes = 0
        for i in range(warm + steps):
            barrier()
            t0 = time.perf_counter()
            tree = tm.make(gids[i % len(gids)])
            dt = time.perf_counter() - t0
            if i >= warm:
                ts = tm.last_stats
                t_sum += dt; cells += ts.hist_cells; rows += ts.hist_rows; rows_idx += ts.hist_rows_indexed
                built += ts.hist_built; hist_ms += ts.gpu_hist_ms; nodes = int(tree.n_nodes)
        barrier()
        v = torch.tensor([t_sum], dtype=torch.float64, device="cuda")
        c = torch.tensor([cells], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(v, op=dist.ReduceOp.MAX)
            dist.all_reduce(c, op=dist.ReduceOp.SUM)
        return dict(t=float(v.item()), cells_all=float(c.item()), cells=cells, rows=rows, rows_idx=rows_idx, built=built,
                    hist_ms=hist_ms, nodes=nodes, steps=steps)

    def entry(name, what, r, alg_bytes, kernel, extra=None):
        ach = alg_bytes / (r["hist_ms"] * 1e-3) / 1e9 if r["hist_ms"] > 0 else 0.0
        e = {"workload": what, "ms_per_tree": r["t"] / r["steps"] * 1e3, "value": r["cells_all"] / r["t"], "unit": UNIT,
             "nodes_per_tree": r["nodes"], "trees_timed": r["steps"],
             "roofline": {"bound": "hbm", "kernel": kernel, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                          "kernel_ms_per_tree": r["hist_ms"] / r["steps"], "algorithmic_bytes_per_tree": alg_bytes / r["steps"]}}
        if extra:
            e.update(extra)
        out[name] = e

    base = dict(min_child_hessian_sum=100.0, min_split_loss=0.0, l1=0.0, l2=0.0, learning_rate=0.1, min_split_samples=-1)
    sampler = ClockSampler(local)
    sampler.mark()
    # ---- configs[2]: 100 M x 128 regression over 8 GPUs = 12.5 M rows per GPU, 255 bins, loss-wise 255 leaves
    n, F, nb = a.c3_rows, 128, 255
    x, gp = _gen_c3(torch, n, F, nb, 42 + rank)
    start = (np.arange(F + 1) * nb).astype(np.int32)
    tm = yb.GpuTreeMaker(yb.make_params(max_depth=-1, max_leaf_cnt=255, grow_policy="loss", **base), start, n, device=local,
                         rank=rank, world=world, nccl_id=fresh_nccl_id())
    for r0 in range(0, n, 1 << 20):
        tm.upload_rows_u8(x[r0:r0 + (1 << 20)], r0)
    tm.set_grads(gp)
    del x
    r = timed(tm, [0], 1, 3)
    tm.close()
    B = int(start[-1])
    entry("c3", "synthetic regression: %d rows/GPU x 128 features, 255 bins, loss-wise 255 leaves (l2 loss)" % n, r,
          r["cells"] + 8 * r["rows"] + 4 * r["rows_idx"] + 16.0 * B * r["built"], "hist_build_k")
    # ---- configs[4]: 20 M x 256, 63 bins, 10 classes, depth 10: class-trees of one boosting round
    n, F, nb, K = a.c5_rows, 256, 63, 10
    x, gp = _gen_c5(torch, n, F, nb, K, 99 + rank)
    start = (np.arange(F + 1) * nb).astype(np.int32)
    tm = yb.GpuTreeMaker(yb.make_params(max_depth=10, max_leaf_cnt=-1, grow_policy="level", class_num=K, **base), start, n,
                         device=local, rank=rank, world=world, nccl_id=fresh_nccl_id())
    for r0 in range(0, n, 1 << 19):
        tm.upload_rows_u8(x[r0:r0 + (1 << 19)], r0)
    tm.set_grads(gp)
    del x, gp
    r = timed(tm, [0, 1, 2, 3], 1, 3)
    tm.close()
    B = int(start[-1])
    entry("c5", "synthetic multiclass: %d rows/GPU x 256 features, 63 bins, 10 classes, level-wise depth 10 "
                "(one tree per class and round; class-trees timed)" % n, r,
          r["cells"] + 8 * r["rows"] + 4 * r["rows_idx"] + 16.0 * B * r["built"], "hist_span_k",
          {"ms_per_round_of_10_class_trees": r["t"] / r["steps"] * 1e3 * K})
    # ---- configs[3]: 50 M x 10 k sparse at 1 % over 8 GPUs = 6.25 M rows per GPU, level-wise depth 6, CSR rows
    n, F = a.c4_rows, 10_000
    start, row_ptr, fids, slots, gp = _gen_c4(torch, n, F, 0.01, 7 + rank)
    nnz = len(fids)
    tm = yb.GpuTreeMaker(yb.make_params(max_depth=6, max_leaf_cnt=-1, grow_policy="level", **base), start, n, device=local,
                         rank=rank, world=world, nccl_id=fresh_nccl_id(), sparse_default=np.zeros(F, np.int32), nnz=nnz)
    step = 1 << 18
    for r0 in range(0, n, step):
        r1 = min(n, r0 + step)
        tm.upload_csr(row_ptr[r0:r1 + 1], fids, slots, row0=r0)
    tm.set_grads(gp)
    del fids, slots
    r = timed(tm, [0], 1, 2)
    tm.close()
    per_row = nnz / max(1, n)
    entry("c4", "synthetic sparse: %d rows/GPU x 10000 features at 1%% (CSR, %d stored entries), up to 255 bins, "
                "level-wise depth 6" % (n, nnz), r, r["rows"] * (per_row * 3 + 20), "hist_csr_k",
          {"bins_total": int(start[-1]), "histogram_mb_per_node": 16.0 * int(start[-1]) / 1e6,
           "value_note": "rows x features counts all 10000 features of a row, stored or not (the dense path's work)"})
    out["clocks"] = sampler.stop()
    return out


def collectives_info(a, world, agg, B):
    """What crosses NVLink per tree and how fast (north_star: the reduce-scatter as a fraction of the NVLink
    roofline).  Every built node's histogram is reduce-scattered by feature range: a rank pulls its 1/N of the
    16*B bytes from each of the other N-1 ranks.  The time is the stream time of the exchange kernels, which
    includes waiting for the slowest rank to arrive -- at this shape the exchanges are latency-, not wire-bound."""
    if world == 1:
        return {"rendezvous_per_tree": 0, "note": "one GPU: no exchange"}
    steps = max(1, a.steps)
    pulled = agg["built"] / steps * 16.0 * B * (world - 1) / world
    ms = agg["ts_comm"] / steps
    ach = pulled / (ms * 1e-3) / 1e9 if ms > 0 else 0.0
    return {"rendezvous_per_tree": agg.get("rendezvous", 0),
            "exchange_ms_per_tree": ms, "split_record_merge_ms_per_tree": agg["ts_sync"] / steps,
            "hist_reduce_scatter": {"bytes_pulled_per_gpu_per_tree": pulled, "achieved_gbs_per_gpu": ach,
                                    "peak_gbs_per_direction": 770.0, "nominal_gbs_per_direction": 900.0,
                                    "frac_of_measured_peer_copy": ach / 770.0,
                                    "peak_source": "B200_PROFILING.md: measured peer copy 770 GB/s per direction (nominal 900)"},
            "note": "latency-bound: %.0f KB per GPU and level at this shape (0.1 us of wire time); the time is "
                    "arrival skew between the ranks at %d rendezvous per tree" % (16.0 * B * (world - 1) / world / 1e3,
                                                                                  agg.get("rendezvous", 0))}


def config(a, world):
    return {"workload": "higgs-shape synthetic: %d rows/GPU x %d features, %d bins, %s-wise depth %d%s"
                        % (a.rows, a.features, a.bins, a.policy, a.depth,
                           "" if a.leaves < 0 else " leaves %d" % a.leaves),
            "rows_per_gpu": a.rows, "global_rows": a.rows * world, "features": a.features, "bins": a.bins,
            "grow_policy": a.policy, "max_depth": a.depth, "max_leaf_cnt": a.leaves,
            "min_child_hessian_sum": 100.0, "l2": 0.0, "learning_rate": 0.1, "loss": "sigmoid",
            "parallelism": "rows sharded x%d, feature-range reduce-scatter" % world,
            "l2_note": "inputs (bins %d MB + grads per GPU) exceed the 126 MB L2, no flush needed"
                       % (a.rows * 32 // 2 ** 20)}


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    xs, ys = synth_bins(min(a.rows, a.cpu_rows), a.features, a.bins, 1234)
    cores = os.cpu_count() or 1
    t0 = time.time()
    for _ in range(a.warmup if a.warmup < 2 else 1):
        cpu_baseline(a, xs, ys, steps=1, threads=cores)
    res = cpu_baseline(a, xs, ys, steps=max(1, a.steps), threads=cores)
    line = {"impl": "reference", "metric": METRIC, "value": res["value"], "unit": UNIT, "n_gpus": a.gpus,
            "steps": a.steps, "warmup": a.warmup, "ms_per_step": res["seconds_per_tree_on_sample"] * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": config(a, a.gpus), "cpu_baseline": res,
            "e2e": {"value": res["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "wall_s": time.time() - t0,
            "note": "reference = CPU restatement of the Java path (no JVM in this image), %d threads" % cores}
    print(json.dumps(line))


def bind_to_gpu_cpus(gpu):
    """Pin this rank to the CPU cores next to its GPU (`nvidia-smi topo -m`, column "CPU Affinity") BEFORE the pinned
    gradient buffer is allocated and first touched, so that the pages of the 88 MB the rank copies every step sit on
    the GPU's own NUMA node.  Returns what was done, for the JSON line."""
    try:
        out = subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True, timeout=20).stdout
        lines = [l for l in out.splitlines() if l.strip()]
        hdr = next(l for l in lines if "CPU Affinity" in l)
        cols = [c.strip() for c in hdr.split("\t")]
        idx = cols.index("CPU Affinity")
        row = next(l for l in lines if l.startswith("GPU%d\t" % gpu) or l.startswith("GPU%d " % gpu))
        cells = [c.strip() for c in row.split("\t")]
        spec = cells[idx]
        cpus = set()
        for part in spec.split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return "nvidia-smi names cores outside this process's set (%s): unchanged" % spec
        os.sched_setaffinity(0, cpus)
        return "cores %s" % spec
    except Exception as e:          # no topology information: nothing is changed
        return "unchanged (%s)" % type(e).__name__


def main():
    a = parse()
    if a.impl == "reference":
        return run_reference(a)

    import torch
    import torch.distributed as dist
    import _pkg
    yb = _pkg.load()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the hot path has no CPU fallback")

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    cpu_binding = bind_to_gpu_cpus(local)
    def fresh_nccl_id():
        """One NCCL unique id per tree-maker clique, made on rank 0 and handed round by the host."""
        if world == 1:
            return None
        idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            idt.copy_(torch.from_numpy(yb.nccl_unique_id()))
        dist.broadcast(idt, 0)
        return idt.cpu().numpy()

    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    nccl_id = fresh_nccl_id()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    n, F = a.rows, a.features
    xs, ys = synth_bins(n, F, a.bins, 1234 + rank, a.uniform_bins)
    start = (np.arange(F + 1) * a.bins).astype(np.int32)
    params = yb.make_params(**tree_kwargs(a))
    tm = yb.GpuTreeMaker(params, start, n, device=local, rank=rank, world=world, nccl_id=nccl_id)
    chunk = 1 << 20
    for r0 in range(0, n, chunk):
        tm.upload_rows(xs[r0:r0 + chunk].astype(np.int32), r0)

    gp_pinned = torch.empty(2 * n, dtype=torch.float32).pin_memory()
    gp = gp_pinned.numpy()
    score = np.zeros(n, np.float32)
    h2d = gp.nbytes
    agg = dict(t_tree=0.0, t_e2e=0.0, cells=0.0, rows=0.0, rows_idx=0.0, built=0, gpu_hist_ms=0.0, launches=0,
               d2h=0, nodes=0, ts_build=0.0, ts_split=0.0, ts_part=0.0, ts_total=0.0, hist_launches=0, ts_comm=0.0,
               ts_init=0.0, ts_sync=0.0, ts_enq=0.0, ts_wait=0.0, ts_gpu=0)
    sampler = ClockSampler(local)                         # started now, used from the first timed step on
    for step in range(a.warmup + a.steps):
        sigmoid_grads(score, ys, gp)                      # the Java side's job, outside the hot path
        timed = step >= a.warmup
        if timed and sampler.t_mark is None:
            sampler.mark()
        barrier()
        t0 = time.perf_counter()
        tm.set_grads_ptr(gp_pinned.data_ptr(), 0, n)      # H2D from pinned memory
        th = time.perf_counter()
        barrier()                                          # ranks finish their copies at different times
        t1 = time.perf_counter()
        tree = tm.make(0)                                  # returns with the tree on the host
        t2 = time.perf_counter()
        ts = tm.last_stats
        pos = tm.get_positions()
        score += tree.node_info[pos]
        if timed:
            agg["t_tree"] += t2 - t1; agg["t_e2e"] += (th - t0) + (t2 - t1)
            agg["cells"] += ts.hist_cells; agg["rows"] += ts.hist_rows; agg["rows_idx"] += ts.hist_rows_indexed
            agg["built"] += ts.hist_built; agg["gpu_hist_ms"] += ts.gpu_hist_ms
            agg["launches"] += ts.kernel_launches; agg["nodes"] = tree.n_nodes
            agg["hist_launches"] += ts.hist_launches
            agg["allocs"] = agg.get("allocs", 0) + ts.device_allocs
            agg["max_batch"] = max(agg.get("max_batch", 0.0), ts.max_batch_ms)
            agg["slow_steps"] = agg.get("slow_steps", []) + [round((t2 - t1) * 1e3, 2)]
            agg["ts_build"] += ts.build_hist; agg["ts_split"] += ts.find_best_split
            agg["ts_part"] += ts.partition; agg["ts_total"] += ts.total; agg["ts_comm"] += ts.communicate_hist
            agg["ts_init"] += ts.init_stats; agg["ts_sync"] += ts.sync_best_split
            agg["ts_enq"] += ts.enqueue_ms; agg["ts_wait"] += ts.host_wait_ms; agg["ts_gpu"] = ts.stage_times_from_gpu
            agg["rendezvous"] = ts.rendezvous; agg["graph"] = ts.graph_replayed
            agg["d2h"] = sum(getattr(tree, nme).nbytes for nme, _ in yb.Tree.FIELDS)
            last_tree = tree
    barrier()
    clocks = sampler.stop()

    # SURVEY.md §8(f) row 1, for information: whole boosting rounds with the gradients computed on the
    # device (score update + loss/gradients + make), nothing but the tree crossing PCIe
    tm.set_targets("sigmoid", ys)
    tm.compute_grads()
    dev_round = 0.0
    for step in range(2 + a.steps):
        barrier()
        t0 = time.perf_counter()
        tm.make(0)
        tm.update_scores(0)
        tm.compute_grads()
        if step >= 2:
            dev_round += time.perf_counter() - t0
    barrier()

    # For information: the reference's own Higgs setting (experiment/higgs/local_gbdt.conf: loss-wise
    # growth, 255 leaves) on the same rows and gradients; BASELINE.json's metric is quoted on depth 8
    loss_wise = None
    used_peer, host_waits = bool(tm.last_stats.peer_collectives), tm.last_stats.batches + 1
    if a.policy == "level" and not a.no_loss_wise:
        tm.close()
        kw = dict(tree_kwargs(a), grow_policy="loss", max_depth=-1, max_leaf_cnt=255)
        tm = yb.GpuTreeMaker(yb.make_params(**kw), start, n, device=local, rank=rank, world=world,
                             nccl_id=fresh_nccl_id())
        for r0 in range(0, n, chunk):
            tm.upload_rows(xs[r0:r0 + chunk].astype(np.int32), r0)
        tm.set_grads_ptr(gp_pinned.data_ptr(), 0, n)
        t_lw = 0.0
        for step in range(2 + a.steps):
            barrier()
            t0 = time.perf_counter()
            tree = tm.make(0)
            if step >= 2:
                t_lw += time.perf_counter() - t0
        barrier()
        loss_wise = {"ms_per_tree": t_lw / a.steps * 1e3, "nodes_per_tree": int(tree.n_nodes),
                     "host_waits_per_tree": tm.last_stats.batches + 1,
                     "what": "loss-wise growth, 255 leaves (the reference's experiment/higgs conf), same rows"}

    # For information at N > 1: the same 11 M rows split over the N GPUs (strong scaling: the latency floor of a level)
    strong = None
    if world > 1 and not a.no_strong:
        tm.close()
        ns = a.rows // world
        xs2, ys2 = synth_bins(ns, F, a.bins, 777 + rank, a.uniform_bins)
        tm = yb.GpuTreeMaker(params, start, ns, device=local, rank=rank, world=world, nccl_id=fresh_nccl_id())
        tm.upload_rows_u8(xs2)
        g2 = np.empty(2 * ns, np.float32)
        sigmoid_grads(np.zeros(ns, np.float32), ys2, g2)
        tm.set_grads(g2)
        t_s = 0.0
        for step in range(3 + a.steps):
            barrier()
            t0 = time.perf_counter()
            tm.make(0)
            if step >= 3:
                t_s += time.perf_counter() - t0
        barrier()
        v = torch.tensor([t_s], dtype=torch.float64, device="cuda")
        dist.all_reduce(v, op=dist.ReduceOp.MAX)
        strong = {"rows_total": ns * world, "rows_per_gpu": ns, "ms_per_tree": float(v.item()) / a.steps * 1e3,
                  "what": "the headline workload's row count split over the %d GPUs (strong scaling), gradients resident" % world}

    pc = None
    if not a.no_parity_check:
        tm.close()
        pc = parity_check(a, yb, torch, dist, world, rank, local, fresh_nccl_id, last_tree)

    oc = None
    if not a.no_other_configs:
        tm.close()
        oc = other_configs(a, yb, torch, dist, world, rank, local, fresh_nccl_id, barrier, measured_peak()[0])

    vals = torch.tensor([agg["t_tree"], agg["t_e2e"], dev_round], dtype=torch.float64, device="cuda")
    sums = torch.tensor([agg["cells"]], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(vals, op=dist.ReduceOp.MAX)
        dist.all_reduce(sums, op=dist.ReduceOp.SUM)
    t_tree, t_e2e, dev_round = vals.tolist()
    cells_all = float(sums.item())

    if rank == 0:
        peak, peak_src = measured_peak()
        # algorithmic bytes of the histogram builds: n*(Fs*1 + 8 + idx) + 16*B per node
        B = int(start[-1])
        alg = agg["cells"] + 8 * agg["rows"] + 4 * agg["rows_idx"] + 16.0 * B * agg["built"]
        ach = alg / (agg["gpu_hist_ms"] * 1e-3) / 1e9 if agg["gpu_hist_ms"] > 0 else 0.0
        line = {
            "metric": METRIC, "value": cells_all / t_tree, "unit": UNIT, "n_gpus": world, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": t_tree / a.steps * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "int64 fixed-point (f32 in, f64-equivalent sums)",
            "data": "synthetic", "config": config(a, world),
            "seconds_per_tree": t_tree / a.steps, "nodes_per_tree": agg["nodes"],
            "e2e": {"value": cells_all / t_e2e, "unit": UNIT, "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": agg["d2h"], "seconds_per_tree": t_e2e / a.steps},
            "device_gradient_round": {"ms_per_round": dev_round / a.steps * 1e3,
                                      "what": "make + score update + sigmoid loss/gradients, all on the device"},
            "loss_wise_255_leaves": loss_wise,
            "gpu_launches": agg["launches"], "peer_memory_collectives": used_peer,
            "time_stats_ms_per_tree": {"build_hist": agg["ts_build"] / a.steps, "find_best_split": agg["ts_split"] / a.steps,
                                       "partition": agg["ts_part"] / a.steps, "total": agg["ts_total"] / a.steps,
                                       "init_stats": agg["ts_init"] / a.steps, "sync_best_split": agg["ts_sync"] / a.steps,
                                       "stage_times_are_cuda_event_times": bool(agg["ts_gpu"]),
                                       "host_enqueue": agg["ts_enq"] / a.steps, "host_wait": agg["ts_wait"] / a.steps,
                                       "collectives_gpu": agg["ts_comm"] / a.steps,
                                       "host_waits": host_waits, "device_allocs_in_timed_steps": agg.get("allocs", 0),
                                       "longest_batch_ms": agg.get("max_batch", 0.0),
                                       "ms_of_each_step": agg.get("slow_steps", [])},
            "roofline": {"bound": "hbm", "kernel": "hist_span_k" if agg["ts_gpu"] else "hist_build_k", "achieved": ach, "peak": peak,
                         "unit": "GB/s", "frac": ach / peak, "traffic": ncu_traffic()[0], "traffic_source": ncu_traffic()[1],
                         "peak_source": peak_src,
                         "kernel_time_source": ("the kernels' own span on the GPU's global timer (launches replayed from a CUDA "
                                                "graph cannot be bracketed by CUDA events; tests check the two agree)"
                                                if agg["ts_gpu"] else "CUDA events on the library's stream"),
                         "algorithmic_bytes_per_launch": alg / max(1, agg["hist_launches"]),
                         "kernel_us_per_launch": agg["gpu_hist_ms"] * 1e3 / max(1, agg["hist_launches"]),
                         "algorithmic_bytes_per_step": alg / a.steps,
                         "kernel_ms_per_step": agg["gpu_hist_ms"] / a.steps,
                         "kernel_share_of_step": agg["gpu_hist_ms"] * 1e-3 / t_tree,
                         "launches_per_step": agg["hist_launches"] / a.steps,
                         "nodes_built_per_step": agg["built"] / a.steps},
            "collectives": collectives_info(a, world, agg, int(start[-1])),
            "strong_11M": strong,
            "cuda_graph_replay": bool(agg.get("graph", 0)),
            "cpu_binding": cpu_binding,
            "clocks": clocks,
            "parity_check": pc,
            "other_configs": oc,
        }
        if not a.no_cpu_baseline and world == 1:
            line["cpu_baseline"] = cpu_baseline(a, xs, ys, steps=1)
        elif not a.no_cpu_baseline:
            line["cpu_baseline"] = None
        print(json.dumps(line))
    tm.close()
    if world > 1:
        dist.destroy_process_group()
    if pc is not None and not pc["ok"]:
        sys.exit(1)


if __name__ == "__main__":
    main()
