"""Full-shape smoke of the other BASELINE configs (C3 reduced to 5M rows) through the estimator API: runs, times, sanity.
GPU box only; not a parity test (those live in tests/)."""
import sys, time
import numpy as np
sys.path.insert(0, ".")
from synapseml_b200.data import synthetic as S
from synapseml_b200.lightgbm.estimators import LightGBMClassifier, LightGBMRegressor, LightGBMRanker, LocalDataFrame

def run(name, est, df, iters):
    t0 = time.time()
    m = est.fit(df)
    dt = time.time() - t0
    pm = est.performance_measures
    print("%s: total %.2f s  phases %s  iterations %d  ms/iter %.2f  model %d bytes" % (
        name, dt, {k: round(v, 3) for k, v in pm.items() if k.endswith("Time")}, iters, pm["trainingIterationsTime"] / iters * 1e3,
        len(m.getNativeModel())), flush=True)
    return m

which = sys.argv[1:] or ["C3", "C4", "C5"]
if "C3" in which:
    X, y = S.dense100(5_000_000)
    df = LocalDataFrame.from_arrays(X, y, num_partitions=8)
    run("C3 5Mx100 L2 255 leaves float", LightGBMRegressor(numIterations=20, numLeaves=255, maxBin=255, samplingMode="fixed", microBatchSize=1 << 15), df, 20)
    run("C3 5Mx100 L2 255 leaves quantized", LightGBMRegressor(numIterations=20, numLeaves=255, maxBin=255, samplingMode="fixed", microBatchSize=1 << 15, useQuantizedGrad=True), df, 20)
    del X, y, df
if "C4" in which:
    X, y, q = S.mslr_shape(2_000_000)
    df = LocalDataFrame.from_arrays(X, y, num_partitions=1, group=q)
    run("C4 2Mx136 lambdarank", LightGBMRanker(numIterations=20, numLeaves=31, maxBin=255, samplingMode="fixed", microBatchSize=1 << 15, groupCol="group"), df, 20)
    del X, y, df
if "C5" in which:
    X, y = S.epsilon_shape(400_000)
    df = LocalDataFrame.from_arrays(X, y, num_partitions=8)
    run("C5 400kx2000 binary float", LightGBMClassifier(numIterations=10, numLeaves=31, maxBin=255, samplingMode="fixed", microBatchSize=1 << 12), df, 10)
    run("C5 400kx2000 binary quantized", LightGBMClassifier(numIterations=10, numLeaves=31, maxBin=255, samplingMode="fixed", microBatchSize=1 << 12, useQuantizedGrad=True), df, 10)
