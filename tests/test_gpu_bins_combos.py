"""Seeded random columns x random binning options (max_bin from 2 to 2000, min_data_in_bin, use_missing / zero_as_missing, categorical
columns, max_bin_by_feature, sample smaller than the data) through K1 and K2 on the GPU against the oracle: mappers, boundaries /
categories and every bin bit-exact."""
import numpy as np
import pytest

from helpers import NativeDataset, oracle_dataset

pytestmark = pytest.mark.gpu


def random_column(rng, n):
    kind = int(rng.integers(0, 10))
    if kind == 0:
        c = rng.standard_normal(n)
    elif kind == 1:
        c = np.where(rng.random(n) < rng.uniform(0.1, 0.9), 0.0, rng.exponential(1.0, n))
    elif kind == 2:
        c = np.where(rng.random(n) < 0.3, 0.0, rng.standard_normal(n) * 3)
    elif kind == 3:
        c = rng.integers(-5, 6, n).astype(np.float64)
    elif kind == 4:
        c = np.round(rng.standard_normal(n), int(rng.integers(0, 3)))          # few to many distinct values with ties
    elif kind == 5:
        c = np.full(n, float(rng.integers(-2, 3)))                              # constant (trivial unless NaNs are added)
    elif kind == 6:
        c = rng.standard_normal(n) * 1e-36                                      # inside the zero threshold
    elif kind == 7:
        c = np.exp(rng.standard_normal(n) * 3)                                  # heavy tail
    elif kind == 8:
        c = -np.abs(rng.standard_normal(n))                                     # negative only
    else:
        c = np.where(rng.random(n) < 0.97, 0.0, 1.0)                            # sparse indicator
    if rng.random() < 0.35:
        c = np.where(rng.random(n) < rng.uniform(0.01, 0.5), np.nan, c)
    return c


def random_categorical(rng, n):
    kind = int(rng.integers(0, 4))
    if kind == 0:
        c = rng.integers(0, int(rng.integers(2, 40)), n).astype(np.float64)
    elif kind == 1:
        c = np.minimum(rng.geometric(0.05, n), 150).astype(np.float64) + float(rng.integers(0, 1000))     # long tail, offset ids
    elif kind == 2:
        c = rng.integers(-3, 8, n).astype(np.float64)                           # negative ids count as missing
    else:
        c = rng.integers(0, 6, n) + rng.random(n) * 0.99                        # fractional values are cut to ints
    if rng.random() < 0.4:
        c = np.where(rng.random(n) < 0.1, np.nan, c)
    return c


@pytest.mark.parametrize("seed", range(150))
def test_random_columns_and_binning_options(native, oracle, seed):
    rng = np.random.Generator(np.random.PCG64(5000 + seed))
    n, f = 8000, 8
    ncat = int(rng.integers(0, 4))
    cat_cols = sorted(rng.choice(f, ncat, replace=False).tolist())
    X = np.stack([random_categorical(rng, n) if j in cat_cols else random_column(rng, n) for j in range(f)], axis=1)
    pick = lambda *xs: xs[int(rng.integers(0, len(xs)))]
    p = ["max_bin=%d" % pick(2, 3, 15, 63, 255, 256, 300, 1000, 2000), "min_data_in_bin=%d" % pick(1, 3, 10, 50),
         "min_data_in_leaf=%d" % pick(1, 20, 200)]
    p += pick([], ["zero_as_missing=true"], ["use_missing=false"], ["feature_pre_filter=false"])
    if cat_cols:
        p.append("categorical_feature=" + ",".join(map(str, cat_cols)))
    if rng.random() < 0.25:
        p.append("max_bin_by_feature=" + ",".join(str(pick(2, 7, 63, 255, 400)) for _ in range(f)))
    params = " ".join(p) + " verbosity=-1"
    sample_rows = pick(8000, 8000, 3000)            # the sample may be smaller than the data: values outside it meet the bins too
    y = np.zeros(n, dtype=np.float32)
    try:
        nd = NativeDataset(X, y, params, sample_rows=sample_rows, micro_batches=[1000, 777])
    except Exception as e:  # noqa: BLE001
        # the one legitimate refusal: a categorical column that needs more than 256 bins
        assert "Categorical feature" in str(e), (params, str(e))
        ref, _, _ = oracle_dataset(oracle, X, y, params, sample_rows=sample_rows)
        assert max(ref.feature_info(j)["num_bin"] for j in cat_cols) > 256, params
        return
    ref, dss, _ = oracle_dataset(oracle, X, y, params, sample_rows=sample_rows)
    assert nd.num_used() == dss[0].num_used_feature(), params
    inner = 0
    for j in range(f):
        a, b = nd.feature_info(j), ref.feature_info(j)
        assert a == b, (params, j, a, b)
        if a["is_trivial"]:
            continue
        assert np.array_equal(nd.bounds(j), ref.bin_upper_bounds(j), equal_nan=True), (params, j)
        assert np.array_equal(nd.bin_column(inner).astype(np.uint16), dss[0].bin_column(inner)), (params, j)
        inner += 1
    nd.free()
