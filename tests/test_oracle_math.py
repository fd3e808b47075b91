import math

import numpy as np


def ulp_diff(a, b):
    return abs(np.float64(a).view(np.int64) - np.float64(b).view(np.int64))


def test_soft_exp_within_one_ulp_of_libm(oracle):
    lib = oracle.lib()
    rng = np.random.Generator(np.random.PCG64(1))
    xs = np.concatenate([rng.uniform(-30, 30, 20000), rng.uniform(-700, 700, 2000), [0.0, 1.0, -1.0, 0.5, -0.5]])
    worst = max(ulp_diff(lib.ORC_SoftExp(float(x)), math.exp(x)) for x in xs)
    assert worst <= 1
    assert lib.ORC_SoftExp(0.0) == 1.0


def test_soft_exp_gradients_match_libm_after_float32_cast(oracle):
    """binary logloss: g = float32(-y / (1 + exp(y s))).  The software exp almost never changes the float32 result."""
    lib = oracle.lib()
    rng = np.random.Generator(np.random.PCG64(2))
    s = rng.uniform(-8, 8, 50000)
    diff = 0
    for v in s:
        a = np.float32(-1.0 / (1.0 + lib.ORC_SoftExp(float(v))))
        b = np.float32(-1.0 / (1.0 + math.exp(v)))
        diff += a != b
    assert diff <= 2     # residue of the <= 1 ulp double difference at float32 rounding boundaries


def test_soft_log2(oracle):
    lib = oracle.lib()
    rng = np.random.Generator(np.random.PCG64(3))
    xs = np.concatenate([np.exp(rng.uniform(-20, 20, 20000)), [1.0, 2.0, 0.5, 1024.0, 3.0]])
    worst = max(ulp_diff(lib.ORC_SoftLog2(float(x)), math.log2(x)) for x in xs if abs(math.log2(x)) > 1e-3)
    assert worst <= 2
    assert lib.ORC_SoftLog2(1.0) == 0.0 and lib.ORC_SoftLog2(8.0) == 3.0


def test_binary_gradients_closed_form(oracle):
    # score 0 (boost_from_average off): response = -y/2 -> g = -/+0.5, h = 0.25
    X = np.arange(40, dtype=np.float64).reshape(40, 1) + 1
    y = (np.arange(40) % 2).astype(np.float32)
    bst, _ = oracle.fit(X, y, "objective=binary boost_from_average=false min_data_in_leaf=1 num_leaves=2 verbosity=-1", num_iterations=1)
    g, h = bst.gradients()
    assert np.array_equal(g, np.where(y > 0, -0.5, 0.5).astype(np.float32))
    assert np.array_equal(h, np.full(40, 0.25, dtype=np.float32))


def test_l2_one_split_known_answer(oracle):
    # x=1 -> y=0 (20 rows), x=2 -> y=1 (20 rows); init = mean = 0.5; g = -/+0.5, h = 1.
    # Best split between 1 and 2: gain = 10^2/20 + 10^2/20 - 0 = 10; leaf outputs -/+0.5; lr = 1; + bias -> leaf values 0 and 1.
    X = np.array([[1.0]] * 20 + [[2.0]] * 20)
    y = np.array([0.0] * 20 + [1.0] * 20, dtype=np.float32)
    bst, _ = oracle.fit(X, y, "objective=regression learning_rate=1.0 min_data_in_leaf=1 min_data_in_bin=1 num_leaves=2 verbosity=-1", num_iterations=1)
    from synapseml_b200.lightgbm.model_text import TextModel
    m = TextModel(bst.save_model_to_string())
    t = m.trees[0]
    assert t.num_leaves == 2 and t.split_feature.tolist() == [0]
    assert abs(t.split_gain[0] - 10.0) < 1e-12
    assert t.leaf_value.tolist() == [0.0, 1.0] and t.leaf_count.tolist() == [20, 20]
    assert 1.0 < t.threshold[0] < 2.0
    assert np.array_equal(bst.scores(), y.astype(np.float64))


def test_quantized_integer_gradients(oracle):
    # deterministic rounding, 4 bins: g/scale in [-2,2] rounded half away from zero, h/scale in [0,4]
    rng = np.random.Generator(np.random.PCG64(5))
    X = rng.standard_normal((500, 3))
    y = (rng.random(500) < 0.4).astype(np.float32)
    bst, _ = oracle.fit(X, y, "objective=binary use_quantized_grad=true stochastic_rounding=false min_data_in_leaf=5 verbosity=-1", num_iterations=1)
    g, h = bst.gradients()
    qg, qh, sc = bst.quant_gradients()
    assert sc[0] == np.abs(g).max() / 2 and sc[1] == np.abs(h).max() / 4
    exp_g = np.trunc(g.astype(np.float64) * (1.0 / sc[0]) + np.where(g >= 0, 0.5, -0.5)).astype(np.int8)
    exp_h = np.trunc(h.astype(np.float64) * (1.0 / sc[1]) + 0.5).astype(np.int8)
    assert np.array_equal(qg, exp_g) and np.array_equal(qh, exp_h)
    assert qg.min() >= -2 and qg.max() <= 2 and qh.min() >= 0 and qh.max() <= 4


def test_regression_objective_gradients_closed_form(oracle):
    """First-iteration gradients of huber / fair / poisson / gamma / tweedie against the formulas of upstream's
    regression_objective.hpp written out in numpy (scores = the automatic initial score: mean label, or its log for the log-link
    objectives).  float32 results may differ by one ulp where numpy's exp and the software exp differ in the last bit."""
    rng = np.random.Generator(np.random.PCG64(3))
    n = 400
    X = rng.normal(size=(n, 2))
    y = rng.poisson(np.exp(0.5 * X[:, 0])).astype(np.float32) + np.float32(0.25)
    w = rng.uniform(0.5, 2.0, n).astype(np.float32)
    mean = float(np.sum(y.astype(np.float64)) / n)

    def check(objective, extra, s0, gfun):
        params = "objective=%s %s min_data_in_leaf=1 num_leaves=2 verbosity=-1" % (objective, extra)
        bst, _ = oracle.fit(X, y, params, num_iterations=1)
        g, h = bst.gradients()
        eg, eh = gfun(np.full(n, s0), y.astype(np.float64))
        assert np.allclose(g, eg.astype(np.float32), rtol=2e-7, atol=0), objective
        assert np.allclose(h, eh.astype(np.float32), rtol=2e-7, atol=0), objective
        assert ("objective=%s\n" % objective) in bst.save_model_to_string()

    a = 0.7
    check("huber", "alpha=%g" % a, mean,
          lambda s, t: (np.where(np.abs(s - t) <= a, s - t, np.sign(s - t) * a), np.ones(n)))
    c = 2.0
    check("fair", "fair_c=%g" % c, mean,
          lambda s, t: (c * (s - t) / (np.abs(s - t) + c), c * c / (np.abs(s - t) + c) ** 2))
    check("poisson", "", np.log(mean),
          lambda s, t: (np.exp(s) - t, np.exp(s) * np.exp(0.7)))
    check("gamma", "", np.log(mean),
          lambda s, t: (1.0 - t * np.exp(-s), t * np.exp(-s)))
    rho = 1.3
    check("tweedie", "tweedie_variance_power=%g" % rho, np.log(mean),
          lambda s, t: (-t * np.exp((1 - rho) * s) + np.exp((2 - rho) * s),
                        -t * (1 - rho) * np.exp((1 - rho) * s) + (2 - rho) * np.exp((2 - rho) * s)))
    # weights multiply both
    bst, _ = oracle.fit(X, y, "objective=poisson min_data_in_leaf=1 num_leaves=2 verbosity=-1", num_iterations=1, weight=w)
    g, h = bst.gradients()
    s0 = np.log(np.sum(y.astype(np.float64) * w) / np.sum(w.astype(np.float64)))
    assert np.allclose(g, ((np.exp(s0) - y) * w).astype(np.float32), rtol=3e-7, atol=1e-7)
    assert np.allclose(h, (np.exp(s0) * np.exp(0.7) * w).astype(np.float32), rtol=3e-7, atol=0)
