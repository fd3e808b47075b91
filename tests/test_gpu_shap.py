"""Feature contributions (predict type 3, featuresShapCol; LightGBMBooster.scala:148,418-427): the GPU TreeSHAP kernel against a
Python restatement of upstream's Tree::TreeSHAP, against brute-force Shapley values on a small model, and the reference's own
property test: the contributions of a row sum to its prediction (VerifyLightGBMRankerStream.scala:73)."""
import itertools
import math

import numpy as np
import pytest

from synapseml_b200.data import synthetic as S
from synapseml_b200.lightgbm.model_text import TextModel

pytestmark = pytest.mark.gpu


class _El:
    __slots__ = ("f", "z", "o", "w")

    def __init__(self, f=-1, z=0.0, o=0.0, w=0.0):
        self.f, self.z, self.o, self.w = f, z, o, w


def _tree_shap(t, x, phi):
    """Upstream Tree::TreeSHAP / ExtendPath / UnwindPath / UnwoundPathSum (src/io/tree.cpp), in Python."""
    def count(nd):
        return float(t.internal_count[nd] if nd >= 0 else t.leaf_count[~nd])

    def decide(nd):
        v = x[t.split_feature[nd]]
        mt = (t.decision_type[nd] >> 2) & 3
        dl = (t.decision_type[nd] & 2) != 0
        isnan = math.isnan(v)
        if isnan and mt != 2:
            v = 0.0
        if (mt == 1 and abs(v) <= 1e-35) or (mt == 2 and isnan):
            return dl
        return v <= t.threshold[nd]

    def extend(path, d, z, o, f):
        path.append(_El(f, z, o, 1.0 if d == 0 else 0.0))
        for i in range(d - 1, -1, -1):
            path[i + 1].w += o * path[i].w * (i + 1) / (d + 1)
            path[i].w = z * path[i].w * (d - i) / (d + 1)

    def unwind(path, d, pi):
        o, z = path[pi].o, path[pi].z
        nxt = path[d].w
        for i in range(d - 1, -1, -1):
            if o != 0:
                tmp = path[i].w
                path[i].w = nxt * (d + 1) / ((i + 1) * o)
                nxt = tmp - path[i].w * z * (d - i) / (d + 1)
            else:
                path[i].w = (path[i].w * (d + 1)) / (z * (d - i))
        for i in range(pi, d):
            path[i].f, path[i].z, path[i].o = path[i + 1].f, path[i + 1].z, path[i + 1].o
        path.pop()

    def unwound_sum(path, d, pi):
        o, z = path[pi].o, path[pi].z
        nxt = path[d].w
        total = 0.0
        for i in range(d - 1, -1, -1):
            if o != 0:
                tmp = nxt * (d + 1) / ((i + 1) * o)
                total += tmp
                nxt = path[i].w - tmp * z * ((d - i) / (d + 1))
            else:
                total += (path[i].w / z) / ((d - i) / (d + 1))
        return total

    def rec(nd, d, parent, pz, po, pf):
        path = [_El(e.f, e.z, e.o, e.w) for e in parent[:d]]
        extend(path, d, pz, po, pf)
        if nd < 0:
            for i in range(1, d + 1):
                w = unwound_sum(path, d, i)
                phi[path[i].f] += w * (path[i].o - path[i].z) * t.leaf_value[~nd]
            return
        left = decide(nd)
        hot = t.left_child[nd] if left else t.right_child[nd]
        cold = t.right_child[nd] if left else t.left_child[nd]
        w = count(nd)
        hz, cz = count(hot) / w, count(cold) / w
        iz = io = 1.0
        pi = next((i for i in range(d + 1) if path[i].f == t.split_feature[nd]), None)
        if pi is not None:
            iz, io = path[pi].z, path[pi].o
            unwind(path, d, pi)
            d -= 1
        rec(hot, d + 1, path, hz * iz, io, t.split_feature[nd])
        rec(cold, d + 1, path, cz * iz, 0.0, t.split_feature[nd])

    total = float(t.internal_count[0]) if t.num_leaves > 1 else 1.0
    phi[-1] += sum(t.leaf_count[i] / total * t.leaf_value[i] for i in range(t.num_leaves)) if t.num_leaves > 1 else t.leaf_value[0]
    if t.num_leaves > 1:
        rec(0, 0, [], 1.0, 1.0, -1)


def _fit(objective="binary", n=30000, iters=6, leaves=15, **kw):
    from synapseml_b200.lightgbm.estimators import LightGBMClassifier, LocalDataFrame
    X, y = S.higgs_shape(n)
    if objective == "multiclass":
        y = np.digitize(X[:, 0] + 0.5 * X[:, 2], [-0.6, 0.7]).astype(np.float32)
    est = LightGBMClassifier(objective=objective, numIterations=iters, numLeaves=leaves, maxBin=63, samplingMode="fixed", microBatchSize=8192, **kw)
    return est.fit(LocalDataFrame.from_arrays(X, y)), X


def test_contributions_equal_the_python_treeshap_and_sum_to_the_raw_score(native):
    model, X = _fit()
    rows = X[:40].copy()
    rows[3, 0] = np.nan
    got = model.predict_contrib(rows)
    assert got.shape == (40, X.shape[1] + 1)
    m = TextModel(model.model_str)
    want = np.zeros_like(got)
    for r in range(len(rows)):
        for t in m.trees:
            _tree_shap(t, rows[r], want[r])
    assert np.allclose(got, want, rtol=1e-10, atol=1e-13)
    assert np.allclose(got.sum(axis=1), model.predict_raw(rows), rtol=0, atol=1e-10)      # the reference's own property


def test_brute_force_shapley_on_a_small_tree(native):
    """Exact Shapley values by enumerating feature subsets, with the tree's own cover counts defining the value of a coalition."""
    model, X = _fit(n=5000, iters=1, leaves=6)
    m = TextModel(model.model_str)
    t = m.trees[0]
    feats = sorted(set(t.split_feature.tolist()))
    x = X[7]

    def value(S_, nd=0):
        if nd < 0:
            return t.leaf_value[~nd]
        f = t.split_feature[nd]
        l, r = t.left_child[nd], t.right_child[nd]
        if f in S_:
            return value(S_, l if x[f] <= t.threshold[nd] else r)
        cl = t.internal_count[l] if l >= 0 else t.leaf_count[~l]
        cr = t.internal_count[r] if r >= 0 else t.leaf_count[~r]
        return (cl * value(S_, l) + cr * value(S_, r)) / (cl + cr)

    M = len(feats)
    want = np.zeros(X.shape[1] + 1)
    for f in feats:
        others = [g for g in feats if g != f]
        for k in range(M):
            for sub in itertools.combinations(others, k):
                wgt = math.factorial(k) * math.factorial(M - k - 1) / math.factorial(M)
                want[f] += wgt * (value(set(sub) | {f}) - value(set(sub)))
    want[-1] = value(set())
    got = model.predict_contrib(x.reshape(1, -1))[0]
    assert np.allclose(got, want, rtol=1e-9, atol=1e-12)


def test_multiclass_and_random_forest_contributions_sum_to_the_raw_scores(native):
    model, X = _fit("multiclass", iters=4)
    c = model.predict_contrib(X[:200]).reshape(200, 3, -1)
    assert np.allclose(c.sum(axis=2), model.predict_raw(X[:200]), rtol=0, atol=1e-10)
    model, X = _fit(iters=5, boostingType="rf", baggingFraction=0.7, baggingFreq=1)
    c = model.predict_contrib(X[:200])
    assert np.allclose(c.sum(axis=1), model.predict_raw(X[:200]), rtol=0, atol=1e-10)
