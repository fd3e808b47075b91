"""Generates the committed golden fixtures from the CPU oracle (run from the repo root: python tests/golden/make_golden.py).

The reference's own implementation of this path (lightgbmlib 3.3.510) cannot be run here, so these vectors pin the oracle
(itself checked against hand-derived answers in tests/test_oracle_bins.py / test_oracle_math.py) and give the GPU tests a
file-based target that does not depend on the oracle being rebuilt identically.
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from helpers import edge_case_matrix, strip_parameters  # noqa: E402
from oracle import oracle as O  # noqa: E402
from synapseml_b200.data import synthetic as S  # noqa: E402


def main():
    n, seed, sample = 6000, 3, 4000
    params = "max_bin=63 min_data_in_leaf=20 bin_construct_sample_cnt=200000"
    X = edge_case_matrix(n, seed)
    ref = O.Dataset.from_sample(X[:sample], 1, n, params)
    ds = ref.create_by_reference(n)
    ds.init_streaming()
    ds.push_rows(X, 0, np.zeros(n, dtype=np.float32))
    ds.mark_finished()
    feats = []
    for j in range(X.shape[1]):
        info = ref.feature_info(j)
        b = ref.bin_upper_bounds(j)
        feats.append(dict(num_bin=info["num_bin"], missing_type=info["missing_type"], default_bin=info["default_bin"],
                          most_freq_bin=info["most_freq_bin"], is_trivial=info["is_trivial"],
                          bounds_hex=["nan" if np.isnan(x) else float.hex(float(x)) for x in b],
                          bounds_hex_as_float=[0.0 if np.isnan(x) or np.isinf(x) else float(x) for x in b]))
    used = ds.num_used_feature()
    import hashlib
    cols = [ds.bin_column(j) for j in range(used)]
    digest = hashlib.sha256(np.stack(cols).astype(np.uint8).tobytes()).hexdigest()
    first_rows = np.stack(cols, axis=1)[:64].astype(int).tolist()
    json.dump(dict(n=n, seed=seed, sample=sample, params=params, features=feats, num_used=used, bins_sha256=digest, first_rows=first_rows),
              open(os.path.join(HERE, "edge_bins.json"), "w"), indent=0)
    X, y = S.higgs_shape(20000)
    pq = ("objective=binary num_leaves=15 learning_rate=0.1 max_bin=63 min_data_in_leaf=20 use_quantized_grad=true "
          "stochastic_rounding=false verbosity=-1")
    bst, _ = O.fit(X, y, pq, num_iterations=5)
    open(os.path.join(HERE, "quant_model.txt"), "w").write(strip_parameters(bst.save_model_to_string()))
    print("golden fixtures written")


if __name__ == "__main__":
    main()
