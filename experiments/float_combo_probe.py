"""One-off probe (not a test: float-mode trees may legitimately differ at near-ties): the option combinations of
tests/test_gpu_combos.py in FLOAT mode, GPU against the oracle — reports crashes, first-tree mismatches and score differences."""
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
bad = []
rels = []
for seed in range(int(sys.argv[1]) if len(sys.argv) > 1 else 100):
    objective, params, weighted = tc.random_combo(seed)
    params = params.replace(" use_quantized_grad=true", "")
    X, z, rng = tc.combo_data(12000, seed)
    y = tc.labels_for(objective, z, rng)
    w = rng.uniform(0.5, 2.0, len(y)).astype(np.float32) if weighted else None
    query = None
    if objective in ("lambdarank", "rank_xendcg"):
        sizes = []
        while sum(sizes) < len(y):
            sizes.append(int(rng.integers(5, 61)))
        query = np.repeat(np.arange(len(sizes)), sizes)[:len(y)].astype(np.int32)
    try:
        nd = NativeDataset(X, y, params, weight=w, query=query, micro_batches=[4096, 100])
        nb = NativeBooster(nd, params)
        _, dss, _ = oracle_dataset(oracle, X, y, params, weight=w, query=query)
        ob = oracle.Booster(dss, params)
        for _ in range(4):
            nb.update(); ob.update_one_iter()
        a, b = TextModel(nb.model()), TextModel(ob.save_model_to_string())
        same_first = a.structure()[:1] == b.structure()[:1]
        same_all = a.structure() == b.structure()
        sa, sb = nb.scores(), ob.scores()
        rel = float(np.abs(sa - sb).max() / max(np.abs(sb).max(), 1e-12))
        rels.append(rel)
        if not same_first or (same_all and rel > 1e-5):
            bad.append((seed, same_first, same_all, rel, params))
        nb.free(); nd.free()
    except Exception as e:  # noqa: BLE001
        bad.append((seed, "EXC", str(e)[:200], params))
print("combos", len(rels), "median rel", float(np.median(rels)), "p90", float(np.percentile(rels, 90)), "max", max(rels))
print("identical structure and rel<=1e-5:", sum(r <= 1e-5 for r in rels))
for b in bad:
    print("BAD", b)
