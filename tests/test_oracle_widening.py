"""Oracle vs hand-derived known answers and independent brute force for the parts of upstream restated in round 2: categorical
bin mapper and split search, features with more than 256 bins, monotone constraints, multiclassova, balanced bagging.  The expected
values are derived in the comments or computed by code that shares nothing with the oracle (numpy brute force), so they do not come
from the code under test.  CPU only."""
import itertools

import numpy as np

from synapseml_b200.lightgbm.model_text import TextModel


def bits_of(words):
    return sorted(32 * w + b for w, x in enumerate(words) for b in range(32) if (int(x) >> b) & 1)


def test_categorical_bin_mapper_known_answer(oracle):
    # 20 sampled rows of one categorical column: 7 x6, 3 x5, 1 x4, 9 x2 (count 2 < min_data_in_bin = 3), NaN x1, -2 x2.
    # Negative values count as missing: na_cnt = 1 + 2 = 3, rest = 17, cut_cnt = round(17 * 0.99) = 17.
    # By falling count: 7 (6), 3 (5), 1 (4) -> used 15 < 17, next is 9 with count 2 < min_data_in_bin and cur_cat = 3 > 1 -> stop.
    # bins: 0 = everything else (-1), 1 = 7, 2 = 3, 3 = 1; not every category got a bin -> missing_type NaN (2).
    # cnt_in_bin = [20 - 15, 6, 5, 4]; default_bin = bin of category 0 = none -> 0; most frequent bin 1 (6/20 < 0.7) -> falls back to 0.
    col = np.array([7.0] * 6 + [3.0] * 5 + [1.0] * 4 + [9.0] * 2 + [np.nan] + [-2.0] * 2)
    X = col.reshape(-1, 1)
    ds = oracle.Dataset.from_sample(X, 1, len(col), "categorical_feature=0 min_data_in_bin=3 min_data_in_leaf=1 max_bin=255 verbosity=-1")
    info = ds.feature_info(0)
    assert (info["num_bin"], info["missing_type"], info["default_bin"], info["most_freq_bin"], info["is_trivial"]) == (4, 2, 0, 0, False)
    assert ds.bin_upper_bounds(0).tolist() == [-1.0, 7.0, 3.0, 1.0]
    d = ds.create_by_reference(8)
    d.init_streaming()
    rows = np.array([[7.0], [3.9], [1.0], [9.0], [np.nan], [-5.0], [0.0], [1e12]])      # 3.9 is cut to 3; 9, NaN, negative, 0, huge -> bin 0
    d.push_rows(rows, 0, np.zeros(8, dtype=np.float32))
    d.mark_finished()
    assert d.bin_column(0).tolist() == [1, 2, 3, 0, 0, 0, 0, 0]
    # ties in the count keep the smaller category first (a stable sort by falling count over ascending values); a category 0 that is
    # seen gets a bin of its own, which is then the default bin
    col = np.array([0.0] * 4 + [5.0] * 4 + [2.0] * 4)
    ds = oracle.Dataset.from_sample(col.reshape(-1, 1), 1, 12, "categorical_feature=0 min_data_in_bin=1 min_data_in_leaf=1 verbosity=-1")
    assert ds.bin_upper_bounds(0).tolist() == [-1.0, 0.0, 2.0, 5.0]
    info = ds.feature_info(0)
    assert (info["missing_type"], info["default_bin"]) == (0, 1)         # every category has a bin -> missing type None


def test_one_vs_rest_categorical_split_known_answer(oracle):
    # three categories (4 bins <= max_cat_to_onehot): 0 -> y = 0 (30 rows), 1 -> y = 0 (30 rows), 2 -> y = 3 (30 rows); L2, init = 1.
    # g = score - y: +1, +1, -2 per row; h = 1.  Category 2 against the rest: gain = 60^2/60 + 60^2/30 = 180, the others:
    # 30^2/30 + 30^2/60 = 45.  Left = the category: output -(-60)/30 = 2 -> leaf 3 after the bias; right -60/60 = -1 -> 0.
    X = np.repeat([0.0, 1.0, 2.0], 30).reshape(-1, 1)
    y = np.repeat([0.0, 0.0, 3.0], 30).astype(np.float32)
    bst, _ = oracle.fit(X, y, "objective=regression learning_rate=1.0 min_data_in_leaf=1 min_data_in_bin=1 num_leaves=2 categorical_feature=0 verbosity=-1",
                        num_iterations=1)
    t = TextModel(bst.save_model_to_string()).trees[0]
    assert t.num_cat == 1 and t.decision_type.tolist() == [1] and t.cat_boundaries.tolist() == [0, 1]
    assert bits_of(t.cat_threshold) == [2]
    assert abs(t.split_gain[0] - 180.0) < 1e-9
    assert np.allclose(t.leaf_value, [3.0, 0.0], atol=1e-12) and t.leaf_count.tolist() == [30, 60]
    assert np.array_equal(bst.scores(), y.astype(np.float64))


def test_many_vs_many_categorical_split_is_the_best_partition(oracle):
    # Sorting the categories by sum_g / sum_h and trying the prefixes from both ends finds the best of ALL 2-partitions of the
    # categories for a squared loss (no smoothing, no extra regularisation, no group-size limits): checked against brute force.
    rng = np.random.Generator(np.random.PCG64(3))
    ncat = 7
    means = rng.normal(0, 2, ncat)
    cats = rng.integers(0, ncat, 700)
    y = (means[cats] + rng.normal(0, 0.3, len(cats))).astype(np.float32)
    X = cats.astype(np.float64).reshape(-1, 1)
    params = ("objective=regression learning_rate=1.0 num_leaves=2 min_data_in_leaf=1 min_data_in_bin=1 min_sum_hessian_in_leaf=0 categorical_feature=0 "
              "max_cat_to_onehot=1 cat_smooth=0 cat_l2=0 min_data_per_group=1 max_cat_threshold=64 verbosity=-1")
    bst, _ = oracle.fit(X, y, params, num_iterations=1)
    t = TextModel(bst.save_model_to_string()).trees[0]
    left = set(bits_of(t.cat_threshold))
    g = (np.float32(np.mean(y.astype(np.float64))) - y).astype(np.float64)      # float32 gradients of the first iteration, h = 1
    def gain(subset):
        m = np.isin(cats, list(subset))
        return g[m].sum() ** 2 / m.sum() + g[~m].sum() ** 2 / (~m).sum() - g.sum() ** 2 / len(g)
    best = max(gain(s) for k in range(1, ncat) for s in itertools.combinations(range(ncat), k))
    assert abs(gain(left) - best) < 1e-6 * best and abs(t.split_gain[0] - best) < 1e-4 * best
    assert 0 < len(left) <= (ncat + 1) // 2                  # the search only looks at prefixes of at most half the categories


def test_wide_bins_follow_searchsorted(oracle):
    # max_bin = 1000 on 5000 distinct values: more than 256 bins; ValueToBin is the first bin whose upper bound covers the value
    rng = np.random.Generator(np.random.PCG64(5))
    x = rng.normal(size=5000)
    ds = oracle.Dataset.from_sample(x.reshape(-1, 1), 1, len(x), "max_bin=1000 min_data_in_bin=1 min_data_in_leaf=1 verbosity=-1")
    info = ds.feature_info(0)
    assert 256 < info["num_bin"] <= 1000
    d = ds.create_by_reference(len(x))
    d.init_streaming()
    d.push_rows(x.reshape(-1, 1), 0, np.zeros(len(x), dtype=np.float32))
    d.mark_finished()
    bounds = ds.bin_upper_bounds(0)
    assert np.array_equal(d.bin_column(0), np.searchsorted(bounds, x, side="left").astype(np.uint16))
    counts = np.bincount(d.bin_column(0), minlength=info["num_bin"])
    assert counts.max() <= 2 * len(x) / info["num_bin"] + 2           # equal-frequency bins


def test_monotone_constraints_known_answer_and_property(oracle):
    # x: 1, 2, 3 (20 rows each), y: 0, 2, 1.  Unconstrained: the tree follows the data (2 -> 2, 3 -> 1).  With an increasing
    # constraint on x the split {1} | {2, 3} is kept (outputs -1 | +0.5 around the mean 1), and the right child may not be split
    # into a higher left and a lower right: one split only.
    X = np.repeat([1.0, 2.0, 3.0], 20).reshape(-1, 1)
    y = np.repeat([0.0, 2.0, 1.0], 20).astype(np.float32)
    base = "objective=regression learning_rate=1.0 min_data_in_leaf=1 min_data_in_bin=1 num_leaves=4 verbosity=-1"
    free, _ = oracle.fit(X, y, base, num_iterations=1)
    assert np.allclose(free.scores(), y)
    mono, _ = oracle.fit(X, y, base + " monotone_constraints=1", num_iterations=1)
    t = TextModel(mono.save_model_to_string()).trees[0]
    assert t.num_leaves == 2 and np.allclose(sorted(t.leaf_value), [0.0, 1.5])
    sc = mono.scores()
    assert np.allclose(sc[:20], 0.0) and np.allclose(sc[20:], 1.5)
    # the property on noisy data, both directions, with a penalty
    rng = np.random.Generator(np.random.PCG64(1))
    Xr = rng.normal(size=(4000, 3))
    yr = (Xr[:, 0] + np.sin(4 * Xr[:, 0]) - Xr[:, 1] + 0.2 * rng.normal(size=4000)).astype(np.float32)
    b, _ = oracle.fit(Xr, yr, "objective=regression num_leaves=15 min_data_in_leaf=10 monotone_constraints=1,-1,0 monotone_penalty=0.7 verbosity=-1",
                      num_iterations=15)
    tm = TextModel(b.save_model_to_string())
    grid = np.linspace(-2.5, 2.5, 40)
    probe = Xr[:100]
    for col, sign in ((0, 1), (1, -1)):
        prev = None
        for v in grid:
            a = probe.copy()
            a[:, col] = v
            p = tm.predict_raw(a)
            assert prev is None or (sign * (p - prev) >= -1e-12).all()
            prev = p


def test_multiclassova_gradients_and_initial_scores_closed_form(oracle):
    # 3 classes with 50 / 30 / 20 rows; per class k a binary problem "label == k": init = log(p / (1 - p)), and at that score
    # g = sigmoid(init) - [label == k] = p - [label == k], h = p (1 - p).
    y = np.repeat([0.0, 1.0, 2.0], [50, 30, 20]).astype(np.float32)
    X = np.random.Generator(np.random.PCG64(0)).normal(size=(100, 2))
    bst, _ = oracle.fit(X, y, "objective=multiclassova num_class=3 num_leaves=2 min_data_in_leaf=1000 verbosity=-1", num_iterations=1)
    g, h = bst.gradients()
    g, h = g.reshape(3, -1), h.reshape(3, -1)
    for k, p in enumerate((0.5, 0.3, 0.2)):
        assert np.allclose(g[k], p - (y == k), atol=1e-6) and np.allclose(h[k], p * (1 - p), atol=1e-6)
    sc = bst.scores().reshape(3, -1)
    for k, p in enumerate((0.5, 0.3, 0.2)):
        assert np.allclose(sc[k], np.log(p / (1 - p)), atol=1e-12)       # no split possible: constant trees carrying the initial scores
    assert "objective=multiclassova num_class:3 sigmoid:1" in bst.save_model_to_string()


def test_balanced_bagging_keeps_the_two_fractions(oracle):
    # pos_bagging_fraction / neg_bagging_fraction: a row is kept when its block's next draw is below the fraction of its label
    rng = np.random.Generator(np.random.PCG64(2))
    X = rng.normal(size=(40000, 3))
    y = (rng.random(40000) < 0.25).astype(np.float32)
    params = ("objective=binary num_leaves=4 bagging_freq=1 pos_bagging_fraction=0.9 neg_bagging_fraction=0.3 bagging_seed=11 learning_rate=0.5 "
              "verbosity=-1")
    bst, _ = oracle.fit(X, y, params, num_iterations=1)
    t = TextModel(bst.save_model_to_string()).trees[0]
    in_bag = t.internal_count[0]                  # rows the root saw
    want = 0.9 * (y > 0).sum() + 0.3 * (y == 0).sum()
    assert abs(in_bag - want) < 4 * np.sqrt(len(y))


def test_cross_entropy_closed_form(oracle):
    # labels are probabilities; init = logit(mean label); at that score g = p - y, h = p (1 - p)
    y = np.array([0.1, 0.9, 0.4, 0.6, 0.25, 0.75] * 20, dtype=np.float32)
    X = np.random.Generator(np.random.PCG64(0)).normal(size=(len(y), 2))
    bst, _ = oracle.fit(X, y, "objective=cross_entropy num_leaves=2 min_data_in_leaf=1000 verbosity=-1", num_iterations=1)
    p = float(np.mean(y.astype(np.float64)))
    assert np.allclose(bst.scores(), np.log(p / (1 - p)), atol=1e-12)
    g, h = bst.gradients()
    assert np.allclose(g, p - y, atol=1e-6) and np.allclose(h, p * (1 - p), atol=1e-6)


def test_rank_xendcg_gradients_against_a_numpy_restatement(oracle):
    # two queries (4 documents, 1 document); scores start at 0, so rho = 1 / 4.  The ground truth is 2^label - u with u drawn from the
    # query's own LCG (x <- 214013 x + 2531011, u = ((x >> 16) & 0x7fff) / 32768, seeded objective_seed + query).
    y = np.array([2.0, 0.0, 1.0, 0.0, 3.0], dtype=np.float32)
    q = np.array([0, 0, 0, 0, 1], dtype=np.int32)
    X = np.random.Generator(np.random.PCG64(0)).normal(size=(5, 2))
    bst, _ = oracle.fit(X, y, "objective=rank_xendcg objective_seed=11 num_leaves=2 min_data_in_leaf=1000 verbosity=-1", num_iterations=1, query=q)
    g, h = bst.gradients()
    x = 11
    us = []
    for _ in range(4):
        x = (214013 * x + 2531011) & 0xFFFFFFFF
        us.append(np.float32(((x >> 16) & 0x7FFF)) / np.float32(32768.0))
    par = np.array([2.0 ** l for l in y[:4]]) - np.array(us, dtype=np.float64)
    rho = np.full(4, 0.25)
    inv = 1.0 / par.sum()
    t1 = -par * inv + rho
    lam = t1.astype(np.float32)
    p1 = t1 / (1 - rho)
    t2 = rho * (p1.sum() - p1)
    lam = lam + t2.astype(np.float32)
    p2 = t2 / (1 - rho)
    lam = lam + (rho * (p2.sum() - p2)).astype(np.float32)
    assert np.allclose(g[:4], lam, rtol=0, atol=1e-7) and np.allclose(h[:4], 0.25 * 0.75, atol=1e-7)
    assert g[4] == 0 and h[4] == 0                           # a query of one document has no gradient
    assert abs(g[:4].sum()) < 1e-6                           # the first-order terms of a query cancel
