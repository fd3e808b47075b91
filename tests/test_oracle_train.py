"""Self-consistency of the oracle's learners (SURVEY.md §8(c) golden vectors (2)) and the committed model fixture."""
import os

import numpy as np

from helpers import strip_parameters
from synapseml_b200.data import synthetic as S
from synapseml_b200.lightgbm.model_text import TextModel

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
PQ = "objective=binary num_leaves=15 learning_rate=0.1 max_bin=63 min_data_in_leaf=20 use_quantized_grad=true stochastic_rounding=false verbosity=-1"


def test_data_parallel_shard_count_does_not_change_quantized_trees(oracle):
    """Integer histograms are order-free: 2, 3 and 5 simulated machines give byte-identical models (data-parallel semantics:
    global estimated counts).  The serial learner (1 machine) uses true counts instead, as upstream does, so it may differ."""
    X, y = S.higgs_shape(30000)
    models = []
    for shards in (2, 3, 5):
        bst, _ = oracle.fit(X, y, PQ, n_shards=shards, num_iterations=8)
        models.append(strip_parameters(bst.save_model_to_string()))
    assert models[0] == models[1] == models[2]


def test_training_scores_equal_model_predictions(oracle):
    X, y = S.higgs_shape(20000)
    for params in (PQ, PQ.replace("use_quantized_grad=true", "use_quantized_grad=false")):
        bst, _ = oracle.fit(X, y, params, num_iterations=10)
        m = TextModel(bst.save_model_to_string())
        assert np.allclose(m.predict_raw(X), bst.scores(), rtol=0, atol=1e-12)
        assert np.array_equal(bst.predict_raw(X[:500]), m.predict_raw(X[:500]))
        for t in m.trees:
            assert t.leaf_count.sum() == len(y) and (t.leaf_count >= 20).all()


def test_determinism(oracle):
    X, y = S.higgs_shape(10000)
    a, _ = oracle.fit(X, y, PQ, num_iterations=5)
    b, _ = oracle.fit(X, y, PQ, num_iterations=5)
    assert a.save_model_to_string() == b.save_model_to_string()


def test_golden_model_fixture(oracle):
    """tests/golden/quant_model.txt was produced by tests/golden/make_golden.py; the GPU test compares the CUDA library with it."""
    X, y = S.higgs_shape(20000)
    bst, _ = oracle.fit(X, y, PQ, num_iterations=5)
    with open(os.path.join(GOLDEN, "quant_model.txt")) as f:
        assert strip_parameters(bst.save_model_to_string()) == f.read()


def test_missing_values_default_direction(oracle):
    """A feature whose NaNs carry the signal: the learner must route the NaN bin by default_left (decision_type bit 1)."""
    rng = np.random.Generator(np.random.PCG64(9))
    n = 4000
    x = rng.standard_normal(n)
    y = (x > 0.3).astype(np.float32)
    nan_rows = rng.random(n) < 0.2
    y[nan_rows] = 1.0
    X = np.stack([np.where(nan_rows, np.nan, x), rng.standard_normal(n)], axis=1)
    bst, _ = oracle.fit(X, y, "objective=binary num_leaves=4 min_data_in_leaf=20 verbosity=-1", num_iterations=3)
    m = TextModel(bst.save_model_to_string())
    t = m.trees[0]
    assert t.split_feature[0] == 0 and ((t.decision_type[0] >> 2) & 3) == 2
    assert np.allclose(m.predict_raw(X), bst.scores(), atol=1e-12)
    p = 1 / (1 + np.exp(-bst.scores()))
    assert ((p > 0.5) == (y > 0)).mean() > 0.97


def test_lambdarank_runs_and_improves_ndcg(oracle):
    X, rel, qid = S.mslr_shape(6000, f=20)
    bst, _ = oracle.fit(X, rel, "objective=lambdarank num_leaves=15 min_data_in_leaf=10 max_position=20 verbosity=-1", num_iterations=10, query=qid)
    s = bst.scores()
    # mean relevance of each query's top-scored document rises well above the query mean
    tops, means = [], []
    for q in np.unique(qid):
        m = qid == q
        if m.sum() < 5:
            continue
        tops.append(rel[m][np.argmax(s[m])])
        means.append(rel[m].mean())
    assert np.mean(tops) > np.mean(means) + 0.3


class _Lcg:
    """Test-side restatement of upstream's utils/random.h, independent of the oracle's C++ (x = 214013 x + 2531011 mod 2^32)."""

    def __init__(self, seed):
        self.x = seed & 0xFFFFFFFF

    def _step(self):
        self.x = (214013 * self.x + 2531011) & 0xFFFFFFFF
        return self.x

    def next_float(self):
        return np.float32((self._step() >> 16) & 0x7FFF) / np.float32(32768.0)

    def next_int(self, lo, hi):
        return (self._step() & 0x7FFFFFFF) % (hi - lo) + lo

    def sample(self, n, k):
        if k == n:
            return list(range(n))
        if k > 1 and k > n / np.log2(k):
            out = []
            for i in range(n):
                if float(self.next_float()) < (k - len(out)) / (n - i):
                    out.append(i)
            return out
        chosen = set()
        for r in range(n - k, n):
            v = self.next_int(0, r)
            chosen.add(r if v in chosen else v)
        return sorted(chosen)


def test_lcg_known_answers():
    r = _Lcg(2)                                        # 214013 * 2 + 2531011 = 2959037 -> >> 16 = 45
    assert r._step() == 2959037 and (2959037 >> 16) & 0x7FFF == 45
    assert abs(float(_Lcg(2).next_float()) - 45 / 32768) < 1e-12


def test_feature_fraction_draws_one_column_sample_per_tree(oracle):
    X, y = S.higgs_shape(20000)
    F = X.shape[1]
    for frac, seed in ((0.6, 2), (0.1, 7)):            # 0.6: sequential selection sampling; 0.1: Floyd's set sampling (K small)
        bst, _ = oracle.fit(X, y, PQ + " feature_fraction=%g feature_fraction_seed=%d" % (frac, seed), num_iterations=6)
        rng = _Lcg(seed)
        k = max(int(F * frac + 0.5), min(2, F))
        for t in TextModel(bst.save_model_to_string()).trees:
            allowed = set(rng.sample(F, k))
            assert len(allowed) == k
            assert set(t.split_feature[:t.num_leaves - 1].tolist()) <= allowed


def test_bagging_bag_size_streams_and_out_of_bag_scores(oracle):
    X, y = S.higgs_shape(20000 + 100)
    n = len(y)
    frac, freq, seed = 0.6, 2, 5
    bst, _ = oracle.fit(X, y, PQ + " bagging_fraction=%g bagging_freq=%d bagging_seed=%d" % (frac, freq, seed), num_iterations=6)
    m = TextModel(bst.save_model_to_string())
    rands = [_Lcg(seed + b) for b in range((n + 1023) // 1024)]
    for it, t in enumerate(m.trees):
        if it % freq == 0:                             # the block streams run on from one re-bagging to the next
            bag = sum(float(rands[i // 1024].next_float()) < frac for i in range(n))
        assert t.internal_count[0] == bag and t.leaf_count.sum() == bag
    assert 0.55 * n < bag < 0.65 * n
    assert np.allclose(m.predict_raw(X), bst.scores(), rtol=0, atol=1e-12)      # every row is scored, in the bag or not


def test_custom_gradients_reproduce_the_builtin_objective(oracle):
    """TrainOneIter(grad, hess): fed with the gradients the binary objective computes itself, the custom path grows the same trees
    (no automatic initial score on either side), and the model text carries no objective line."""
    from helpers import oracle_dataset
    X, y = S.higgs_shape(15000)
    base = PQ + " boost_from_average=false"
    _, dss, _ = oracle_dataset(oracle, X, y, base)
    a = oracle.Booster(dss, base)
    b = oracle.Booster(dss, base.replace("objective=binary", "objective=custom"))
    for _ in range(5):
        assert not a.update_one_iter()
        g, h = a.gradients()
        assert not b.update_one_iter_custom(g, h)
    ma, mb = a.save_model_to_string(), b.save_model_to_string()
    assert ma.split("Tree=0\n", 1)[1].split("end of trees")[0] == mb.split("Tree=0\n", 1)[1].split("end of trees")[0]
    assert np.array_equal(a.scores(), b.scores())
    assert "objective=binary" in ma and "\nobjective=" not in mb.split("Tree=0")[0]
