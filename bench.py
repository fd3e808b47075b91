#!/usr/bin/env python
"""Benchmark of the LightGBM data-parallel training hot path on B200.

    python bench.py --gpus N --steps K --warmup W            (N > 1: launched by torchrun, one rank per GPU)
    python bench.py --impl reference --gpus N --steps K --warmup W

A step is one boosting iteration (one LGBM_BoosterUpdateOneIter: gradients, quantisation, one leaf-wise tree, score
update) over the whole synthetic dataset.  Workload = BASELINE.json configs[1] (C2): LightGBMClassifier, binary logloss,
Higgs-shape 10M x 28 float64, maxBin=255, numLeaves=31, useQuantizedGrad.  With N GPUs the 10M rows are sharded
contiguously over the ranks (strong scaling, as the metric is defined on a fixed 10M x 28 dataset).

Prints ONE JSON line (rank 0).  `value` = rows * iterations / second of the boosting loop with the binned dataset
resident in HBM; `e2e` = the same metric through the estimator API from host numpy buffers (bin-boundary search, H2D
copies, row binning, the K iterations and the model-string read-back all inside the timed region).
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

N_ROWS = 10_000_000
N_FEAT = 28
BIN_SAMPLE = 200_000
WORKLOAD = ("C2: LightGBMClassifier binary logloss, synthetic Higgs-shape 10Mx28 f64, maxBin=255, numLeaves=31, "
            "learningRate=0.1, minDataInLeaf=20, useQuantizedGrad=true (num_grad_quant_bins=4, stochastic_rounding=true)")
CPU_SAMPLE_ROWS = 2_000_000


def estimator(num_iterations, micro_batch=1 << 16):
    from synapseml_b200.lightgbm.estimators import LightGBMClassifier
    return LightGBMClassifier(numIterations=num_iterations, numLeaves=31, maxBin=255, learningRate=0.1, minDataInLeaf=20,
                              useQuantizedGrad=True, samplingMode="fixed", microBatchSize=micro_batch, verbosity=-1)


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


def cpu_baseline(steps, warmup, rows=CPU_SAMPLE_ROWS):
    """Times the CPU oracle (a from-scratch restatement of upstream LightGBM, not lightgbmlib) on the host cores."""
    from oracle import oracle as O
    from synapseml_b200.data import synthetic as S
    X, y = S.higgs_shape(rows)
    params = ("objective=binary num_leaves=31 learning_rate=0.1 max_bin=255 min_data_in_leaf=20 use_quantized_grad=true "
              "num_grad_quant_bins=4 stochastic_rounding=true verbosity=-1")
    threads = O.max_threads()
    bst, dss = O.fit(X, y, params, num_iterations=None, sample_rows=BIN_SAMPLE)
    for _ in range(warmup):
        bst.update_one_iter()
    t0 = time.time()
    for _ in range(steps):
        bst.update_one_iter()
    dt = time.time() - t0
    bst.free()
    for d in dss:
        d.free()
    return dict(value=rows * steps / dt, unit="rows*iters/s", cores=threads, kind="port",
                sample="first %d rows of the C2 workload, %d iterations after %d warm-up, OpenMP %d threads" % (rows, steps, warmup, threads)), dt / steps * 1e3


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cb, ms = cpu_baseline(args.steps, args.warmup)
    out = {"impl": "reference", "metric": "lightgbm_train_rows_iters_per_sec", "value": cb["value"], "unit": "rows*iters/s",
           "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
           "scaling": "strong", "vs_baseline": None, "dtype": "int32", "data": "synthetic",
           "config": {"workload": WORKLOAD, "note": "reference JVM path (SynapseML + lightgbmlib 3.3.510 under local[N] Spark) is "
                      "unavailable on this box (no JVM); this arm times the CPU oracle (a restatement of upstream LightGBM, not "
                      "lightgbmlib) on a bounded sample"},
           "cpu_baseline": cb,
           "e2e": {"value": cb["value"], "unit": "rows*iters/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(out), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=500)      # C2 trains 500 iterations
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--rows", type=int, default=N_ROWS)
    # strong (default): the named 10M-row dataset is sharded over the GPUs; weak: every GPU gets --rows rows (information only)
    ap.add_argument("--scaling", choices=["strong", "weak"], default="strong")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    args.warmup = max(args.warmup, 3)

    import torch
    import torch.distributed as dist
    from synapseml_b200.data import synthetic as S
    from synapseml_b200.lightgbm import network_manager as nm
    from synapseml_b200.lightgbm.booster import LightGBMBooster
    from synapseml_b200.lightgbm.dataset import LightGBMDataset, ReferenceDatasetUtils
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
    N = args.rows * (world if args.scaling == "weak" else 1)
    lo, hi = N * rank // world, N * (rank + 1) // world
    X, y = S.higgs_shape(hi - lo, start=lo)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    # ------------------------------------------------------------------ resident-in-HBM measurement
    est = estimator(args.steps + args.warmup)
    train_params = est.get_train_params(world, 1)
    ds_params = dataset_creation_params(255, BIN_SAMPLE, 20, train_params.executionParams.numThreads)
    payload = [None]
    if rank == 0:
        payload[0] = ReferenceDatasetUtils.create_reference_dataset_from_sample(ds_params, "features", N, N_FEAT, X[:BIN_SAMPLE])
    if world > 1:
        dist.broadcast_object_list(payload, src=0)
    ctx = dict(train_params=train_params, dataset_params=ds_params, num_cols=N_FEAT, serialized_reference=payload[0])
    task = StreamingPartitionTask(ctx)
    t0 = time.time()
    dataset = task.prepare_dataset([{"features": X, "label": y}])
    ingest_s = time.time() - t0
    info = nm.get_global_network_info() if world > 1 else None
    if info is not None:
        nm.init_lightgbm_network(info)
    booster = LightGBMBooster(dataset, train_params.to_string())
    for _ in range(args.warmup):
        booster.update_one_iteration()
    sa = booster.stats()
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
    t = torch.tensor([dev_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    value = N * args.steps / (ms_total * 1e-3)
    launches = sb["kernel_launches"] - sa["kernel_launches"]
    growth = {k: (sb[k] - sa[k]) / args.steps for k in ("growth_rounds", "growth_expansions", "growth_splits")}
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
    hist_rows = s1["hist_rows_visited"] - s0["hist_rows_visited"]
    hist_ms = s1["hist_ms"] - s0["hist_ms"]
    n_local = hi - lo
    root_rows = n_local * args.steps
    nonroot_rows = max(hist_rows - root_rows, 0)
    hist_launches = s1["hist_launches"] - s0["hist_launches"]
    hist_kernel_launches = max(s1["hist_kernel_launches"] - s0["hist_kernel_launches"], 1)
    bytes_alg = root_rows * (N_FEAT + 2) + nonroot_rows * (N_FEAT + 2 + 4) + hist_launches * N_FEAT * 256 * 16
    peak, peak_src = measured_peak()
    achieved = bytes_alg / (hist_ms * 1e-3) / 1e9 if hist_ms > 0 else 0.0
    root_ms = C.c_float()
    validate(lightgbmlib().LGBM_B200_BoosterTimeRootHistogram(booster.booster_handle, 5, C.byref(root_ms)))
    root_gbs = n_local * (N_FEAT + 2) / (root_ms.value * 1e-3) / 1e9
    traffic, traffic_src = None, None
    tpath = os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "r01_hist_traffic.json")
    if os.path.exists(tpath):       # DRAM bytes per launch from the committed ncu --set full capture of the same workload
        tj = json.load(open(tpath))
        traffic, traffic_src = tj["dram_bytes_per_launch"], tj["source"]
    roofline = {"bound": "hbm", "kernel": "k_bt_hist (histogram construction, K4)", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                "algorithmic_bytes_per_launch": bytes_alg / hist_kernel_launches,
                "avg_launch_us": hist_ms * 1e3 / hist_kernel_launches, "kernel_launches_per_step": hist_kernel_launches / args.steps,
                "peak_source": peak_src,
                "frac_of_spec_8000": achieved / 8000.0,
                "algorithmic_bytes": "per row visit: 28 bin bytes + 2 (int8 g,h) + 4 row index (0 at the root); + F*256*16 output per launch",
                "hist_ms_per_step": hist_ms / args.steps, "hist_share_of_step": hist_ms / ms_timed_pass,
                "rows_visited_per_step": hist_rows / args.steps, "hist_launches_per_step": hist_launches / args.steps,
                "avg_us_per_leaf_histogram": hist_ms * 1e3 / max(hist_launches, 1),
                "root_launch": {"ms": root_ms.value, "achieved": root_gbs, "frac": root_gbs / peak}}
    booster.free_native_memory()
    dataset.close()
    if info is not None:
        nm.free_lightgbm_network()

    # ------------------------------------------------------------------ end to end through the estimator API (host buffers)
    e2e = None
    if not args.no_e2e:
        est2 = estimator(args.steps)
        df = LocalDataFrame.from_arrays(X, y, num_partitions=8)     # 8 task threads push concurrently, as 8 Spark tasks would
        barrier()
        t0 = time.time()
        model = est2.fit(df)
        barrier()
        dt = time.time() - t0
        tt = torch.tensor([dt], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dt = float(tt.item())
        h2d = X.nbytes + y.nbytes + (BIN_SAMPLE * N_FEAT * 8 if rank == 0 else 0)
        e2e = {"value": N * args.steps / dt, "unit": "rows*iters/s", "h2d_bytes_per_step": h2d / args.steps,
               "d2h_bytes_per_step": len(model.model_str) / args.steps, "seconds": dt,
               "phases_seconds": {k: round(float(v), 4) for k, v in (est2.performance_measures or {}).items()
                                  if k.endswith("Time")},
               "includes": "bin-boundary search on the 200k-row sample, H2D of the f64 rows (pinned staging), row binning, "
                           "%d boosting iterations, model string read-back" % args.steps}
    if rank != 0:
        return
    cb = None
    if world == 1 and not args.no_cpu_baseline:
        cb, _ = cpu_baseline(steps=5, warmup=1)
    out = {"metric": "lightgbm_train_rows_iters_per_sec", "value": value, "unit": "rows*iters/s", "n_gpus": world, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": args.scaling,
           "vs_baseline": None, "dtype": "int32", "data": "synthetic",
           "config": {"workload": WORKLOAD, "rows": N, "features": N_FEAT, "rows_per_gpu": n_local, "parallelism": "dp%d" % world,
                      "l2": "inputs larger than L2 (binned matrix %d MB per GPU + 2x%d MB index buffers vs 126 MB L2)" % (n_local * 32 >> 20, n_local * 4 >> 20),
                      "timing": "CUDA events on the library's launching stream, max over ranks; wall %.1f ms" % wall_ms,
                      "ingest_seconds_untimed": ingest_s},
           "clocks": clocks_summary(clk_path, local_rank), "e2e": e2e, "gpu_launches": int(launches), "growth_per_step": growth,
           "roofline": roofline,
           "cpu_baseline": cb}
    print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
