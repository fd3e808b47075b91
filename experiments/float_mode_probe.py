"""C2 shape in float mode (C1's mode at C2's size): per-iteration time and phase profile.  GPU box only."""
import sys, time
import numpy as np
sys.path.insert(0, ".")
sys.path.insert(0, "tests")
from helpers import NativeBooster, NativeDataset
from synapseml_b200.data import synthetic as S
X, y = S.higgs_shape(10_000_000)
for quant in ("false", "true"):
    params = "objective=binary num_leaves=31 max_bin=255 min_data_in_leaf=20 verbosity=-1 use_quantized_grad=%s" % quant
    nd = NativeDataset(X, y, params, micro_batches=[1 << 16])
    nb = NativeBooster(nd, params)
    for _ in range(3): nb.update()
    nb.scores()
    t0 = time.time()
    for _ in range(30): nb.update()
    nb.scores()
    print("use_quantized_grad=%s: %.3f ms per iteration" % (quant, (time.time() - t0) / 30 * 1e3), flush=True)
    nb.free(); nd.free()
