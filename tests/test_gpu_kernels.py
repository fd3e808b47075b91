"""Gradient (K3) and histogram (K4) kernels against the oracle / numpy on the same seeded inputs: bit-exact integers."""
import numpy as np
import pytest

from helpers import NativeBooster, NativeDataset, oracle_dataset
from synapseml_b200.data import synthetic as S

pytestmark = pytest.mark.gpu


def setup(oracle, n, f, params, weight=None, sample_rows=200000):
    X, y = S.higgs_shape(n, f=f)
    nd = NativeDataset(X, y, params, weight=weight, micro_batches=[8192])
    nb = NativeBooster(nd, params)
    _, dss, _ = oracle_dataset(oracle, X, y, params, weight=weight)
    ob = oracle.Booster(dss, params)
    return X, y, nd, nb, dss[0], ob


@pytest.mark.parametrize("extra", ["", "is_unbalance=true", "sigmoid=0.7 scale_pos_weight=1.5"])
def test_binary_gradients_bit_exact_over_iterations(native, oracle, extra):
    params = "objective=binary num_leaves=15 min_data_in_leaf=20 max_bin=63 verbosity=-1 " + extra
    rng = np.random.Generator(np.random.PCG64(1))
    w = rng.uniform(0.5, 2.0, 30000).astype(np.float32) if "sigmoid" in extra else None
    _, _, nd, nb, _, ob = setup(oracle, 30000, 28, params, weight=w)
    for _ in range(3):      # gradients depend on scores, which depend on the (float-mode) trees: compare before each update
        nb.update()
        ob.update_one_iter()
        g, h = nb.gradients()
        og, oh = ob.gradients()
        # float-mode trees agree to ~1e-7; the first iteration's gradients (before any tree) must be bit-exact
        assert np.allclose(g, og, rtol=1e-5, atol=1e-7) and np.allclose(h, oh, rtol=1e-5, atol=1e-7)
    nb.free()
    nd.free()


def test_first_iteration_gradients_and_quantisation_bit_exact(native, oracle):
    for rounding in ("false", "true"):
        params = ("objective=binary num_leaves=15 min_data_in_leaf=20 max_bin=63 use_quantized_grad=true seed=11 "
                  "stochastic_rounding=%s verbosity=-1" % rounding)
        _, _, nd, nb, _, ob = setup(oracle, 50000, 28, params)
        for _ in range(4):   # quantized mode is bit-exact end to end, so every iteration's gradients must match exactly
            nb.update()
            ob.update_one_iter()
            g, h = nb.gradients()
            og, oh = ob.gradients()
            assert np.array_equal(g, og) and np.array_equal(h, oh)
            qg, qh, sc = nb.quant_gradients()
            oqg, oqh, osc = ob.quant_gradients()
            assert np.array_equal(sc, osc) and np.array_equal(qg, oqg) and np.array_equal(qh, oqh)
        nb.free()
        nd.free()


@pytest.mark.parametrize("n,f", [(70000, 28), (40000, 50), (20000, 130), (33, 28), (5000, 3)])
def test_quantized_root_histogram_equals_numpy(native, oracle, n, f):
    """All three row-record layouts (R = 4, 2, 1), ragged sizes, more than one flush per CTA at 70k rows."""
    params = "objective=binary num_leaves=4 min_data_in_leaf=1 max_bin=255 use_quantized_grad=true stochastic_rounding=false verbosity=-1"
    _, _, nd, nb, ods, ob = setup(oracle, n, f, params)
    nb.update()
    qg, qh, _ = nb.quant_gradients()
    hist = nb.root_histogram()
    for j in range(nd.num_used()):
        col = nd.bin_column(j)
        eg = np.bincount(col, weights=qg.astype(np.float64), minlength=256).astype(np.int64)
        eh = np.bincount(col, weights=qh.astype(np.float64), minlength=256).astype(np.int64)
        assert np.array_equal(hist[j, :, 0], eg) and np.array_equal(hist[j, :, 1], eh), j
    # size-independent property: every feature's bins sum to the same leaf totals
    assert (hist[:, :, 0].sum(axis=1) == qg.astype(np.int64).sum()).all() and (hist[:, :, 1].sum(axis=1) == qh.astype(np.int64).sum()).all()
    nb.free()
    nd.free()


@pytest.mark.parametrize("n,f", [(60000, 28), (30000, 50), (9000, 130)])
def test_float_mode_root_histogram_is_exact_fixed_point(native, oracle, n, f):
    """Float mode accumulates 19-bit fixed point with power-of-two scales: sums are exact integers, independent of order."""
    params = "objective=binary num_leaves=4 min_data_in_leaf=1 max_bin=255 verbosity=-1"
    _, _, nd, nb, _, _ = setup(oracle, n, f, params)
    nb.update()
    g, h = nb.gradients()
    eg_scale = 2.0 ** (19 - np.frexp(np.abs(g).max())[1])
    eh_scale = 2.0 ** (19 - np.frexp(np.abs(h).max())[1])
    gi = np.rint(g.astype(np.float64) * eg_scale).astype(np.int64)
    hi = np.rint(h.astype(np.float64) * eh_scale).astype(np.int64)
    hist = nb.root_histogram()
    for j in range(nd.num_used()):
        col = nd.bin_column(j)
        eg = np.zeros(256, dtype=np.int64)
        eh = np.zeros(256, dtype=np.int64)
        np.add.at(eg, col, gi)
        np.add.at(eh, col, hi)
        assert np.array_equal(hist[j, :, 0], eg) and np.array_equal(hist[j, :, 1], eh), j
    nb.free()
    nd.free()
