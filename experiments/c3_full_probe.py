"""C3 at its full single-GPU size (20M x 100, 255 leaves, L2): does it run, how long per iteration, is the fit sane.  GPU box only."""
import sys, time
import numpy as np
sys.path.insert(0, ".")
from synapseml_b200.data import synthetic as S
from synapseml_b200.lightgbm.estimators import LightGBMRegressor, LocalDataFrame
X, y = S.dense100(20_000_000)
df = LocalDataFrame.from_arrays(X, y, num_partitions=8)
for quant in (True, False):
    est = LightGBMRegressor(numIterations=10, numLeaves=255, maxBin=255, samplingMode="fixed", microBatchSize=1 << 15, useQuantizedGrad=quant,
                            isProvideTrainingMetric=True, metric="l2")
    t0 = time.time()
    m = est.fit(df)
    pm = est.performance_measures
    l2 = [h[0]["l2"] for h in pm["evalHistory"]]
    print("quantized=%s total %.1f s  ms/iter %.2f  train l2 %s" % (quant, time.time() - t0, pm["trainingIterationsTime"] / 10 * 1e3,
                                                                   ["%.4f" % v for v in l2[::3]]), flush=True)
    assert all(b < a for a, b in zip(l2, l2[1:]))
pred = m.predict_raw(X[:100000])
print("rmse on the first 100k rows after 10 iterations: %.4f (label std %.4f)" % (np.sqrt(np.mean((pred - y[:100000]) ** 2)), y[:100000].std()))
