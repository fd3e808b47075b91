#!/usr/bin/env python
"""Benchmark of the LightGBM data-parallel training hot path on B200.

    python bench.py --gpus N --steps K --warmup W [--config C1|C2|C3|C4|C5]      (N > 1: launched by torchrun, one rank per GPU)
    python bench.py --impl reference --gpus N --steps K --warmup W [--config ...]

A step is one boosting iteration (one LGBM_BoosterUpdateOneIter: gradients, quantisation / fixed point, one leaf-wise tree, score
update) over the whole synthetic dataset.  The default workload is BASELINE.json configs[1] (C2: LightGBMClassifier, binary
logloss, Higgs-shape 10M x 28 float64, maxBin=255, numLeaves=31, useQuantizedGrad); --config selects one of the other four
(SURVEY.md §8(d) gives their shapes and parameters).  With N GPUs the rows are sharded contiguously over the ranks (strong
scaling: the metric is defined on a fixed dataset).

Prints ONE JSON line (rank 0).  `value` = rows * iterations / second of the boosting loop with the binned dataset resident in
HBM; `e2e` = the same metric through the estimator API from host numpy buffers (bin-boundary search, H2D copies, row binning,
the K iterations and the model-string read-back all inside the timed region); `parity_check` = a small fixed case trained through
the same booster path on the same ranks right before the timed loop and compared, on rank 0, with the CPU oracle's learner for
the same number of machines (a mismatch ends the run with a non-zero exit code).
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

BIN_SAMPLE = 200_000
METRIC = "lightgbm_train_rows_iters_per_sec"

# SURVEY.md §8(d).  gh = bytes of the gradient pair a histogram row visit reads (2: int8 x 2, 8: two f32, 4: f32 + constant hessian)
CONFIGS = {
    "C1": dict(est="classifier", data="higgs", rows=1_000_000, features=28, iterations=100, gh=8,
               kw=dict(numLeaves=31, maxBin=255, learningRate=0.1, minDataInLeaf=20, minSumHessianInLeaf=1e-3),
               text="C1: LightGBMClassifier binary logloss, synthetic Higgs-shape 1Mx28 f64, 100 iterations, numLeaves=31, maxBin=255, float gradients"),
    "C2": dict(est="classifier", data="higgs", rows=10_000_000, features=28, iterations=500, gh=2,
               kw=dict(numLeaves=31, maxBin=255, learningRate=0.1, minDataInLeaf=20, useQuantizedGrad=True),
               text="C2: LightGBMClassifier binary logloss, synthetic Higgs-shape 10Mx28 f64, maxBin=255, numLeaves=31, learningRate=0.1, "
                    "minDataInLeaf=20, useQuantizedGrad=true (num_grad_quant_bins=4, stochastic_rounding=true)"),
    "C3": dict(est="regressor", data="dense100", rows=20_000_000, features=100, iterations=100, gh=4,
               kw=dict(numLeaves=255, maxBin=255, learningRate=0.1, minDataInLeaf=20),
               text="C3: LightGBMRegressor L2, synthetic dense 20Mx100 f64, numLeaves=255, streaming dataTransferMode, float gradients (constant hessian)"),
    "C4": dict(est="ranker", data="mslr", rows=2_000_000, features=136, iterations=100, gh=8,
               kw=dict(numLeaves=31, maxBin=255, learningRate=0.1, minDataInLeaf=20, maxPosition=20, evalAt=[1, 2, 3, 4, 5]),
               text="C4: LightGBMRanker lambdarank, synthetic MSLR-shape 2Mx136 f64 with a query group column, maxPosition=20, float gradients"),
    "C5": dict(est="classifier", data="epsilon", rows=400_000, features=2000, iterations=100, gh=8,
               kw=dict(numLeaves=31, maxBin=255, learningRate=0.1, minDataInLeaf=20),
               text="C5: LightGBMClassifier binary logloss, synthetic Epsilon-shape 400kx2000 f64 (uint8 bins, feature-heavy histograms), float gradients"),
}
CPU_ROW_ITERS_BUDGET = 3.0e8        # bounds the CPU arm: rows * (steps + warmup) (about 15-30 s on 16 host cores for C2)


def estimator(cfg, num_iterations, micro_batch=1 << 16, **extra):
    from synapseml_b200.lightgbm import estimators as E
    cls = {"classifier": E.LightGBMClassifier, "regressor": E.LightGBMRegressor, "ranker": E.LightGBMRanker}[cfg["est"]]
    return cls(numIterations=num_iterations, samplingMode="fixed", microBatchSize=micro_batch, verbosity=-1, **cfg["kw"], **extra)


def make_data(cfg, rows, start=0):
    """(X, y, group) of global rows [start, start + rows) of the config's synthetic dataset."""
    from synapseml_b200.data import synthetic as S
    if cfg["data"] == "higgs":
        X, y = S.higgs_shape(rows, start=start, f=cfg["features"])
        return X, y, None
    if cfg["data"] == "dense100":
        X, y = S.dense100(rows, start=start, f=cfg["features"])
        return X, y, None
    if cfg["data"] == "epsilon":
        X, y = S.epsilon_shape(rows, start=start, f=cfg["features"])
        return X, y, None
    raise ValueError(cfg["data"])


def shard(cfg, total_rows, rank, world):
    """This rank's contiguous row shard (query-aligned for the ranker, LightGBMRanker.scala:88-120) and the bin sample (global
    first rows).  Returns (X, y, group, lo, hi, total_rows, sample)."""
    if cfg["data"] == "mslr":
        from synapseml_b200.data import synthetic as S
        X, y, q = S.mslr_shape(total_rows, f=cfg["features"])
        n = len(y)
        qb = np.concatenate([[0], np.flatnonzero(np.diff(q) != 0) + 1, [n]])
        cut = [0] + [int(qb[np.searchsorted(qb, n * r // world)]) for r in range(1, world)] + [n]
        lo, hi = cut[rank], cut[rank + 1]
        return X[lo:hi], y[lo:hi], q[lo:hi], lo, hi, n, X[:BIN_SAMPLE]
    lo, hi = total_rows * rank // world, total_rows * (rank + 1) // world
    X, y, _ = make_data(cfg, hi - lo, lo)
    sample = X[:BIN_SAMPLE] if rank == 0 else None
    if rank == 0 and hi - lo < min(BIN_SAMPLE, total_rows):
        sample = make_data(cfg, min(BIN_SAMPLE, total_rows), 0)[0]
    return X, y, None, lo, hi, total_rows, sample


def clocks_sampler_start(path):
    q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    try:
        return subprocess.Popen(["nvidia-smi", "--query-gpu=" + q, "--format=csv,noheader,nounits", "-lms", "100"],
                                stdout=open(path, "w"), stderr=subprocess.DEVNULL)
    except OSError:
        return None


def clocks_summary(path, device_index):
    sm, mx, reasons = [], [], set()
    try:
        for line in open(path):
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9 or f[0] != str(device_index):
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
    except OSError:
        pass
    if not sm:
        return {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
    return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons)}


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:  # noqa: BLE001
        return 6650.0, "fallback (B200_PROFILING.md)"


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def cpu_baseline(cfg, steps, warmup, budget=CPU_ROW_ITERS_BUDGET):
    """Times the CPU oracle (a from-scratch restatement of upstream LightGBM, not lightgbmlib) on ALL host cores, whatever
    OMP_NUM_THREADS says (torchrun exports 1), on the config's own rows and parameters: the full dataset when
    rows * (steps + warmup) fits the budget, else its first rows.  One protocol for the in-line cpu_baseline and --impl reference:
    `warmup` untimed iterations, then `steps` timed ones."""
    from oracle import oracle as O
    threads = host_cores()
    O.set_num_threads(threads)
    rows = int(min(cfg["rows"], max(BIN_SAMPLE, budget // (steps + warmup))))
    if cfg["data"] == "mslr":
        X, y, q, _, _, rows, _ = shard(cfg, rows, 0, 1)
    else:
        X, y, q = make_data(cfg, rows)
    params = estimator(cfg, steps + warmup).get_train_params(1, 1).to_string()
    bst, dss = O.fit(X, y, params, num_iterations=None, sample_rows=BIN_SAMPLE, query=q)
    for _ in range(warmup):
        bst.update_one_iter()
    t0 = time.time()
    for _ in range(steps):
        bst.update_one_iter()
    dt = time.time() - t0
    bst.free()
    for d in dss:
        d.free()
    return dict(value=rows * steps / dt, unit="rows*iters/s", cores=threads, kind="port", rows=rows,
                sample="%s of the %s workload (%d of %d rows), %d iterations after %d warm-up, OpenMP %d threads" %
                       ("all rows" if rows == cfg["rows"] else "first rows", cfg["name"], rows, cfg["rows"], steps, warmup, threads)), dt / steps * 1e3


def run_reference(args, cfg):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cb, ms = cpu_baseline(cfg, args.steps, max(args.warmup, 1), budget=8 * CPU_ROW_ITERS_BUDGET)
    out = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": "rows*iters/s",
           "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
           "scaling": "strong", "vs_baseline": None, "dtype": "int32" if cfg["gh"] == 2 else "f64", "data": "synthetic",
           "config": {"workload": cfg["text"], "rows": cb["rows"], "features": cfg["features"],
                      "note": "reference JVM path (SynapseML + lightgbmlib 3.3.510 under local[N] Spark) is unavailable on this box (no JVM); "
                              "this arm times the CPU oracle (a restatement of upstream LightGBM, not lightgbmlib) on all host cores"},
           "cpu_baseline": cb,
           "e2e": {"value": cb["value"], "unit": "rows*iters/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(out), flush=True)


def parity_check(world, rank, info, dist):
    """200k x 28, 8 quantized iterations through the booster path the timed loop uses (same ranks, same exchange) against the CPU
    oracle's learner with `world` machines: model text must be identical."""
    from synapseml_b200.data import synthetic as S
    from synapseml_b200.lightgbm.booster import LightGBMBooster
    from synapseml_b200.lightgbm.dataset import ReferenceDatasetUtils
    from synapseml_b200.lightgbm.estimators import LightGBMClassifier
    from synapseml_b200.lightgbm.params import dataset_creation_params
    from synapseml_b200.lightgbm.partition_task import StreamingPartitionTask
    n, iters, sample = 200_000, 8, 50_000
    lo, hi = n * rank // world, n * (rank + 1) // world
    X, y = S.higgs_shape(hi - lo, start=lo)
    est = LightGBMClassifier(numIterations=iters, numLeaves=31, maxBin=63, useQuantizedGrad=True, stochasticRounding=False,
                             samplingMode="fixed", microBatchSize=8192, binSampleCount=sample, verbosity=-1)
    tp = est.get_train_params(world, 1)
    dsp = dataset_creation_params(63, sample, 20, tp.executionParams.numThreads)
    payload = [None]
    if rank == 0:
        Xs = X[:sample] if hi - lo >= sample else S.higgs_shape(sample)[0]
        payload[0] = ReferenceDatasetUtils.create_reference_dataset_from_sample(dsp, "features", n, 28, Xs)
    if world > 1:
        dist.broadcast_object_list(payload, src=0)
    task = StreamingPartitionTask(dict(train_params=tp, dataset_params=dsp, num_cols=28, serialized_reference=payload[0]))
    ds = task.prepare_dataset([{"features": X, "label": y}])
    bst = LightGBMBooster(ds, tp.to_string())
    for _ in range(iters):
        bst.update_one_iteration()
    ours = bst.save_to_string(None).split("parameters:")[0]
    exchange = bst.get_profile()["exchange"]
    bst.free_native_memory()
    ds.close()
    res = {"n_shards": world, "rows": n, "iterations": iters, "exchange": exchange, "identical": None}
    if rank == 0:
        from oracle import oracle as O
        O.set_num_threads(host_cores())
        Xa, ya = S.higgs_shape(n)
        ob, odss = O.fit(Xa, ya, tp.to_string(), dataset_params="max_bin=63 bin_construct_sample_cnt=%d min_data_in_leaf=20" % sample,
                         n_shards=world, num_iterations=iters, sample_rows=sample)
        res["identical"] = bool(ours == ob.save_model_to_string().split("parameters:")[0])
        ob.free()
        for d in odss:
            d.free()
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None)      # default: the config's own iteration count
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--config", default="C2", choices=sorted(CONFIGS))
    ap.add_argument("--rows", type=int, default=None)
    # strong (default): the named dataset is sharded over the GPUs; weak: every GPU gets --rows rows (information only)
    ap.add_argument("--scaling", choices=["strong", "weak"], default="strong")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity-check", action="store_true")
    ap.add_argument("--no-profile", action="store_true")
    args = ap.parse_args()
    cfg = dict(CONFIGS[args.config], name=args.config)
    if args.rows:
        cfg["rows"] = args.rows
    if args.steps is None:
        args.steps = cfg["iterations"]
    if args.impl == "reference":
        return run_reference(args, cfg)
    args.warmup = max(args.warmup, 3)

    import torch
    import torch.distributed as dist
    from synapseml_b200.lightgbm import network_manager as nm
    from synapseml_b200.lightgbm.booster import LightGBMBooster
    from synapseml_b200.lightgbm.dataset import ReferenceDatasetUtils
    from synapseml_b200.lightgbm.estimators import LocalDataFrame
    from synapseml_b200.lightgbm.params import dataset_creation_params
    from synapseml_b200.lightgbm.partition_task import StreamingPartitionTask
    from synapseml_b200.lightgbm.native import lightgbmlib, validate

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: this library has no CPU path")
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group(backend="cpu:gloo,cuda:nccl", rank=rank, world_size=world)
    F = cfg["features"]
    X, y, group, lo, hi, N, sample = shard(cfg, cfg["rows"] * (world if args.scaling == "weak" else 1), rank, world)
    n_local = hi - lo

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def max_over_ranks(v):
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    info = nm.get_global_network_info() if world > 1 else None
    if info is not None:
        nm.init_lightgbm_network(info)
    pc = None if args.no_parity_check else parity_check(world, rank, info, dist)

    # ------------------------------------------------------------------ resident-in-HBM measurement
    est = estimator(cfg, args.steps + args.warmup)
    train_params = est.get_train_params(world, 1)
    ds_params = dataset_creation_params(255, BIN_SAMPLE, 20, train_params.executionParams.numThreads)
    payload = [None]
    if rank == 0:
        payload[0] = ReferenceDatasetUtils.create_reference_dataset_from_sample(ds_params, "features", N, F, sample)
    if world > 1:
        dist.broadcast_object_list(payload, src=0)
    ctx = dict(train_params=train_params, dataset_params=ds_params, num_cols=F, serialized_reference=payload[0])
    task = StreamingPartitionTask(ctx)
    part = {"features": X, "label": y}
    if group is not None:
        part["group"] = group
    barrier()
    t0 = time.time()
    dataset = task.prepare_dataset([part])
    ingest_s = max_over_ranks(time.time() - t0)
    # K2 alone: the production binning kernel over resident raw rows (the host->device copy is outside)
    k2_rows = int(min(n_local, max(1, (1 << 28) // (8 * F))))
    k2_ms = C.c_float()
    xs = np.ascontiguousarray(X[:k2_rows], dtype=np.float64)
    validate(lightgbmlib().LGBM_B200_DatasetTimeBinning(dataset.dataset_ptr, xs.ctypes.data_as(C.POINTER(C.c_double)), k2_rows, 5, C.byref(k2_ms)))
    del xs
    booster = LightGBMBooster(dataset, train_params.to_string())
    for _ in range(args.warmup):
        booster.update_one_iteration()
    sa = booster.stats()
    pa = booster.get_profile()
    clk_path = os.path.join(tempfile.gettempdir(), "bench_clocks_%d.csv" % rank)
    sampler = clocks_sampler_start(clk_path) if rank == 0 else None
    barrier()
    booster.timer_start()
    w0 = time.time()
    for _ in range(args.steps):
        booster.update_one_iteration()
    dev_ms = booster.timer_stop()
    barrier()
    wall_ms = (time.time() - w0) * 1e3
    if sampler is not None:
        sampler.terminate()
    sb = booster.stats()
    ms_total = max_over_ranks(dev_ms)
    value = N * args.steps / (ms_total * 1e-3)
    launches = sb["kernel_launches"] - sa["kernel_launches"]
    growth = {k: (sb[k] - sa[k]) / args.steps for k in ("growth_rounds", "growth_expansions", "growth_splits")}
    pb = booster.get_profile()
    growth.update({k: (pb[k] - pa[k]) / args.steps for k in ("rounds_enqueued", "tree_refetches")})
    # second pass of the same length with every histogram launch bracketed by CUDA events on the launching stream (the events
    # break the programmatic overlap between launches, so this pass is a little slower): roofline numbers only
    booster.set_hist_timing(True)
    s0 = booster.stats()
    booster.timer_start()
    for _ in range(args.steps):
        booster.update_one_iteration()
    ms_timed_pass = booster.timer_stop()
    s1 = booster.stats()
    booster.set_hist_timing(False)

    # roofline of the histogram kernel over the timed region (rank 0's launches)
    gh = cfg["gh"]
    hist_rows = s1["hist_rows_visited"] - s0["hist_rows_visited"]
    hist_ms = s1["hist_ms"] - s0["hist_ms"]
    root_rows = n_local * args.steps
    nonroot_rows = max(hist_rows - root_rows, 0)
    hist_launches = s1["hist_launches"] - s0["hist_launches"]
    hist_kernel_launches = max(s1["hist_kernel_launches"] - s0["hist_kernel_launches"], 1)
    bytes_alg = root_rows * (F + gh) + nonroot_rows * (F + gh + 4) + hist_launches * F * 256 * 16
    peak, peak_src = measured_peak()
    achieved = bytes_alg / (hist_ms * 1e-3) / 1e9 if hist_ms > 0 else 0.0
    root_ms = C.c_float()
    validate(lightgbmlib().LGBM_B200_BoosterTimeRootHistogram(booster.booster_handle, 5, C.byref(root_ms)))
    root_gbs = n_local * (F + gh) / (root_ms.value * 1e-3) / 1e9
    traffic = traffic_src = line_bound = None
    tpath = os.path.join(ROOT, "profiles", "r02_hist_traffic.json")
    if not os.path.exists(tpath):
        tpath = os.path.join(ROOT, "profiles", "r01_hist_traffic.json")
    if args.config == "C2" and os.path.exists(tpath):       # DRAM bytes per launch from the committed ncu --set full capture of this workload
        tj = json.load(open(tpath))
        traffic, traffic_src = tj["dram_bytes_per_launch"], tj["source"]
        line_bound = tj.get("line_granular_note")
    roofline = {"bound": "hbm", "kernel": "k_bt_hist (histogram construction, K4)", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src, "line_granular_bound": line_bound,
                "algorithmic_bytes_per_launch": bytes_alg / hist_kernel_launches,
                "avg_launch_us": hist_ms * 1e3 / hist_kernel_launches, "kernel_launches_per_step": hist_kernel_launches / args.steps,
                "peak_source": peak_src,
                "frac_of_spec_8000": achieved / 8000.0,
                "algorithmic_bytes": "per row visit: %d bin bytes + %d (gradient pair) + 4 row index (0 at the root); + F*256*16 output per leaf histogram" % (F, gh),
                "hist_ms_per_step": hist_ms / args.steps, "hist_share_of_step": hist_ms / ms_timed_pass,
                "rows_visited_per_step": hist_rows / args.steps, "hist_launches_per_step": hist_launches / args.steps,
                "avg_us_per_leaf_histogram": hist_ms * 1e3 / max(hist_launches, 1),
                "root_launch": {"ms": root_ms.value, "achieved": root_gbs, "frac": root_gbs / peak}}

    # the other HBM-bound kernels (SURVEY.md §8(d) algorithmic bytes), timed per kernel type with CUDA events in a short third pass
    roofline_other = {"K2_row_binning": {"kernel": "k2_bin_rows", "rows": k2_rows, "ms": k2_ms.value,
                                         "achieved": k2_rows * F * 9 / (k2_ms.value * 1e-3) / 1e9, "unit": "GB/s",
                                         "algorithmic_bytes": "N*F*(8 read + 1 written)"}}
    roofline_other["K2_row_binning"]["frac"] = roofline_other["K2_row_binning"]["achieved"] / peak
    if not args.no_profile:
        psteps = min(args.steps, 10)
        booster.set_profile(True)
        p0 = booster.get_profile()
        for _ in range(psteps):
            booster.update_one_iteration()
        p1 = booster.get_profile()
        booster.set_profile(False)

        def kms(name):
            a, b = p0["kernels"].get(name, {"ms": 0.0}), p1["kernels"].get(name, {"ms": 0.0})
            return b["ms"] - a["ms"]
        part_rows = p1["partition_rows"] - p0["partition_rows"]
        fused = cfg["est"] != "ranker"          # binary / L2: the apply-tree pass also produces the next gradients (K8 + K3 in one pass)
        rows_p = n_local * psteps
        specs = [("K3_quantise", "quantize", rows_p * (8 + (2 if gh == 2 else 8)), "N*(8 read f32 g,h + %d written)" % (2 if gh == 2 else 8)),
                 ("K7_partition", "partition", part_rows * (4 + 1 + 4), "rows partitioned * (4 index read + 1 bin + 4 index written)"),
                 ("K8_score_update" + ("_plus_K3_gradients" if fused else ""), "score_update", rows_p * (F + 8 + 8 + ((4 + 8) if fused else 0)),
                  "N*(F bins + 8 score read + 8 written" + (" + 4 label + 8 next g,h written)" if fused else ")"))]
        if not fused:
            specs.append(("K3_gradients", "gradients", rows_p * (8 + 4 + 8), "N*(8 score + 4 label + 8 g,h written); lambdarank is pair-bound, not HBM-bound"))
        for key, name, nbytes, note in specs:
            ms = kms(name)
            if ms > 0:
                roofline_other[key] = {"ms_per_step": ms / psteps, "achieved": nbytes / (ms * 1e-3) / 1e9, "unit": "GB/s",
                                       "frac": nbytes / (ms * 1e-3) / 1e9 / peak, "algorithmic_bytes": note}
        roofline_other["per_kernel_ms_per_step"] = {k: round((p1["kernels"][k]["ms"] - p0["kernels"].get(k, {"ms": 0.0})["ms"]) / psteps, 5)
                                                    for k in p1["kernels"]}
        roofline_other["note"] = "third pass of %d iterations with CUDA events around every launch (no overlap between launches)" % psteps
    exchange = booster.get_profile()["exchange"]
    booster.free_native_memory()
    dataset.close()

    # ------------------------------------------------------------------ end to end through the estimator API (host buffers)
    e2e = None
    if not args.no_e2e:
        if info is not None:      # the estimator brings its own network up, as a fit in the reference does per task
            nm.free_lightgbm_network()
            info = None
        est2 = estimator(cfg, args.steps)
        df = LocalDataFrame.from_arrays(X, y, num_partitions=8, group=group) if group is None else LocalDataFrame([part])
        barrier()
        t0 = time.time()
        model = est2.fit(df)
        barrier()
        dt = max_over_ranks(time.time() - t0)
        h2d = X.nbytes + y.nbytes + (BIN_SAMPLE * F * 8 if rank == 0 else 0)
        pm = est2.performance_measures or {}
        e2e = {"value": N * args.steps / dt, "unit": "rows*iters/s", "h2d_bytes_per_step": h2d / args.steps,
               "d2h_bytes_per_step": len(model.model_str) / args.steps, "seconds": dt,
               "phases_seconds": {k: round(float(v), 4) for k, v in pm.items() if k.endswith("Time")},
               "booster_setup_ms": pm.get("boosterSetupMs"),
               "includes": "bin-boundary search on the 200k-row sample, H2D of the f64 rows (pinned staging), row binning, network + "
                           "booster set-up, %d boosting iterations, model string read-back (rank 0's phases; seconds = max over ranks)" % args.steps}
    if info is not None:
        nm.free_lightgbm_network()
    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return
    cb = None
    if world == 1 and not args.no_cpu_baseline:
        cb, _ = cpu_baseline(cfg, args.steps, args.warmup)
    out = {"metric": METRIC, "value": value, "unit": "rows*iters/s", "n_gpus": world, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": args.scaling,
           "vs_baseline": None, "dtype": "int32" if gh == 2 else "f64", "data": "synthetic",
           "config": {"workload": cfg["text"], "name": args.config, "rows": N, "features": F, "rows_per_gpu": n_local, "parallelism": "dp%d" % world,
                      "exchange": exchange,
                      "l2": "inputs larger than L2 (binned matrix %d MB per GPU + 2x%d MB index buffers vs 126 MB L2)" % (n_local * max(32, F) >> 20, n_local * 4 >> 20),
                      "timing": "CUDA events on the library's launching stream, max over ranks; wall %.1f ms" % wall_ms,
                      "ingest_seconds_untimed": ingest_s,
                      "ingest_gbs": n_local * world * F * 8 / ingest_s / 1e9},
           "clocks": clocks_summary(clk_path, local_rank), "e2e": e2e, "gpu_launches": int(launches), "growth_per_step": growth,
           "parity_check": pc, "roofline": roofline, "roofline_other": roofline_other,
           "cpu_baseline": cb}
    print(json.dumps(out), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if pc is not None and pc["identical"] is False:
        raise SystemExit("parity_check failed: the model trained on %d rank(s) differs from the oracle's" % world)


if __name__ == "__main__":
    main()
