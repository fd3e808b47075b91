import sys, os
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
import numpy as np
from helpers import NativeDataset, NativeBooster
from synapseml_b200.data import synthetic as S
X, y = S.higgs_shape(2000000)
params = "objective=binary num_leaves=31 max_bin=255 use_quantized_grad=true seed=4242 verbosity=-1"
nd = NativeDataset(X, y, params, micro_batches=[1<<16]); nb = NativeBooster(nd, params)
for _ in range(12): nb.update()
nb.free(); nd.free()
