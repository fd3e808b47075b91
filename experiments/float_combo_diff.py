"""One-off: for given seeds of the float-mode combination probe, show the first node where the GPU's and the oracle's first tree
differ, with both gains (a near-tie is legitimate in float mode; a large gap is not)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from helpers import NativeBooster, NativeDataset, oracle_dataset  # noqa: E402
from oracle import oracle  # noqa: E402
from synapseml_b200.lightgbm.model_text import TextModel  # noqa: E402
import test_gpu_combos as tc  # noqa: E402

oracle.build()
for seed in map(int, sys.argv[1:]):
    objective, params, weighted = tc.random_combo(seed)
    params = params.replace(" use_quantized_grad=true", "")
    X, z, rng = tc.combo_data(12000, seed)
    y = tc.labels_for(objective, z, rng)
    w = rng.uniform(0.5, 2.0, len(y)).astype(np.float32) if weighted else None
    nd = NativeDataset(X, y, params, weight=w, micro_batches=[4096, 100])
    nb = NativeBooster(nd, params)
    _, dss, _ = oracle_dataset(oracle, X, y, params, weight=w)
    ob = oracle.Booster(dss, params)
    nb.update(); ob.update_one_iter()
    a, b = TextModel(nb.model()).trees[0], TextModel(ob.save_model_to_string()).trees[0]
    print("seed", seed, params)
    print(" leaves", a.num_leaves, b.num_leaves)
    for i in range(min(a.num_leaves, b.num_leaves) - 1):
        same = a.split_feature[i] == b.split_feature[i] and a.threshold[i] == b.threshold[i]
        if not same:
            print("  node", i, "gpu: f", a.split_feature[i], "thr", a.threshold[i], "gain", a.split_gain[i], "cnt", a.internal_count[i],
                  "| oracle: f", b.split_feature[i], "thr", b.threshold[i], "gain", b.split_gain[i], "cnt", b.internal_count[i])
            break
    else:
        print("  identical up to the shorter tree")
    print("  gains gpu   ", np.round(a.split_gain[:8], 4).tolist())
    print("  gains oracle", np.round(b.split_gain[:8], 4).tolist())
    nb.free(); nd.free()
