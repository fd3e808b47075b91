"""Parity at the shapes BASELINE.json names, not at reduced ones: the benchmarked configuration C2 (10M x 28, quantized gradients,
stochastic rounding on and off) and C1 (1M x 28, float mode, 100 iterations) trained through the C ABI and by the oracle on the same
rows; the row-binning kernel (K2) against the oracle's bin matrix at 1M x 28 and at the Epsilon shape 400k x 2000 (SURVEY.md §7.1,
§8(c)).  Quantized: byte-identical model text and training scores.  Float: identical first trees, raw scores within 1e-5 relative
(north_star's tolerance)."""
import hashlib

import numpy as np
import pytest

from helpers import NativeBooster, NativeDataset, oracle_dataset, strip_parameters
from synapseml_b200.data import synthetic as S
from synapseml_b200.lightgbm.model_text import TextModel

pytestmark = pytest.mark.gpu
FLOAT_RTOL = 1e-5


def test_c2_as_benchmarked_quantized_model_is_byte_identical(native, oracle):
    """bench.py's workload, row for row: 10M x 28 Higgs-shape, 31 leaves, max_bin 255, 4 gradient bins; 20 iterations with
    stochastic rounding (the throughput run) and without (the parity run of SURVEY.md §8(d))."""
    X, y = S.higgs_shape(10_000_000)
    base = ("objective=binary num_leaves=31 learning_rate=0.1 max_bin=255 min_data_in_leaf=20 use_quantized_grad=true "
            "num_grad_quant_bins=4 verbosity=-1")
    nd = NativeDataset(X, y, base, micro_batches=[1 << 18])
    _, dss, _ = oracle_dataset(oracle, X, y, base)
    for sr in ("true", "false"):
        params = base + " stochastic_rounding=" + sr
        nb = NativeBooster(nd, params)
        ob = oracle.Booster(dss, params)
        for _ in range(20):
            assert not nb.update()
            assert not ob.update_one_iter()
        assert strip_parameters(nb.model()) == strip_parameters(ob.save_model_to_string()), "stochastic_rounding=" + sr
        assert np.array_equal(nb.scores(), ob.scores())
        nb.free()
        ob.free()
    nd.free()
    for d in dss:
        d.free()


def test_c1_float_mode_100_iterations(native, oracle):
    """C1: 1M x 28, float gradients, 100 iterations.  The oracle accumulates float32 gradients into float64 bins as upstream does;
    the library accumulates fixed-point integers — the first trees must be the same trees and the raw scores of
    the whole 100-iteration model must stay within north_star's 1e-5."""
    X, y = S.higgs_shape(1_000_000)
    params = "objective=binary num_leaves=31 learning_rate=0.1 max_bin=255 min_data_in_leaf=20 min_sum_hessian_in_leaf=1e-3 verbosity=-1"
    nd = NativeDataset(X, y, params, micro_batches=[1 << 18])
    nb = NativeBooster(nd, params)
    _, dss, _ = oracle_dataset(oracle, X, y, params)
    ob = oracle.Booster(dss, params)
    a10 = b10 = None
    for it in range(100):
        assert not nb.update()
        assert not ob.update_one_iter()
        if it == 9:
            a10, b10 = nb.scores(), ob.scores()
    sn, so = TextModel(nb.model()).structure(), TextModel(ob.save_model_to_string()).structure()
    assert sn[:5] == so[:5]                                    # identical split features, thresholds and topology
    assert np.abs(a10 - b10).max() <= FLOAT_RTOL * np.abs(b10).max()
    a, b = nb.scores(), ob.scores()
    assert np.abs(a - b).max() <= FLOAT_RTOL * np.abs(b).max(), np.abs(a - b).max() / np.abs(b).max()
    nb.free()
    ob.free()
    nd.free()
    for d in dss:
        d.free()


def _bins_digest(cols):
    h = hashlib.sha256()
    for c in cols:
        h.update(np.ascontiguousarray(c, dtype=np.uint8).tobytes())
    return h.hexdigest()


@pytest.mark.parametrize("shape", ["higgs_1m_x_28", "epsilon_400k_x_2000"])
def test_k2_bin_matrix_sha256_at_full_shape(native, oracle, shape):
    if shape == "higgs_1m_x_28":
        X, y = S.higgs_shape(1_000_000)
    else:
        X, y = S.epsilon_shape(400_000)
    params = "max_bin=255 bin_construct_sample_cnt=200000 min_data_in_leaf=20 verbosity=-1"
    nd = NativeDataset(X, y, params, micro_batches=[1 << 16])
    _, dss, _ = oracle_dataset(oracle, X, y, params)
    nu = nd.num_used()
    assert nu == dss[0].num_used_feature() == X.shape[1]
    ours = _bins_digest(nd.bin_column(j) for j in range(nu))
    theirs = _bins_digest(dss[0].bin_column(j) for j in range(nu))
    assert ours == theirs
    # the row-major copy the histogram kernel reads holds the same bins
    rm = nd.row_major(0, 4096)
    for j in (0, nu // 2, nu - 1):
        assert np.array_equal(rm[:, j], nd.bin_column(j)[:4096])
    nd.free()
    for d in dss:
        d.free()
