"""Reader for the LightGBM text model with a vectorised numpy predictor.

Used by the model classes for feature importances and by the tests, as an independent traversal, to compare models and scores
produced by the CUDA library and by the oracle.  (Scoring itself goes through the library: LGBM_BoosterPredictForMat on the GPU.)
"""
import numpy as np

K_ZERO_THRESHOLD = float(np.float32(1e-35))


class TextTree:
    def __init__(self, kv):
        self.num_leaves = int(kv["num_leaves"])
        self.shrinkage = float(kv.get("shrinkage", 1))
        f = lambda k, t: np.array(kv[k].split(), dtype=t) if k in kv and kv[k] else np.zeros(0, dtype=t)
        self.leaf_value = f("leaf_value", np.float64)
        if self.num_leaves > 1:
            self.split_feature = f("split_feature", np.int64)
            self.split_gain = f("split_gain", np.float64)
            self.threshold = f("threshold", np.float64)
            self.decision_type = f("decision_type", np.int64)
            self.left_child = f("left_child", np.int64)
            self.right_child = f("right_child", np.int64)
            self.leaf_weight = f("leaf_weight", np.float64)
            self.leaf_count = f("leaf_count", np.int64)
            self.internal_value = f("internal_value", np.float64)
            self.internal_weight = f("internal_weight", np.float64)
            self.internal_count = f("internal_count", np.int64)
            # categorical splits (decision_type bit 0): threshold = index k of the node's category set, words
            # [cat_boundaries[k], cat_boundaries[k + 1]) of cat_threshold, one bit per category value
            self.num_cat = int(kv.get("num_cat", 0))
            self.cat_boundaries = f("cat_boundaries", np.int64)
            self.cat_threshold = f("cat_threshold", np.uint64)

    def predict_leaf(self, X):
        n = X.shape[0]
        if self.num_leaves <= 1:
            return np.zeros(n, dtype=np.int64)
        node = np.zeros(n, dtype=np.int64)
        active = np.ones(n, dtype=bool)
        while active.any():
            idx = np.flatnonzero(active)
            nd = node[idx]
            fval = X[idx, self.split_feature[nd]]
            mt = (self.decision_type[nd] >> 2) & 3
            default_left = (self.decision_type[nd] & 2) != 0
            isnan = np.isnan(fval)
            fval = np.where(isnan & (mt != 2), 0.0, fval)
            is_missing = ((mt == 1) & (np.abs(fval) <= K_ZERO_THRESHOLD)) | ((mt == 2) & isnan)
            go_left = np.where(is_missing, default_left, fval <= self.threshold[nd])
            is_cat = (self.decision_type[nd] & 1) != 0
            if is_cat.any():      # Tree::CategoricalDecision: NaN, negative and unlisted values go right
                raw = X[idx, self.split_feature[nd]]
                ok = is_cat & ~np.isnan(raw) & (raw < 2147483648.0)
                iv = np.where(ok, raw, 0.0).astype(np.int64)
                ok &= iv >= 0
                k = np.where(is_cat, self.threshold[nd], 0).astype(np.int64)
                lo, hi = self.cat_boundaries[k], self.cat_boundaries[np.minimum(k + 1, len(self.cat_boundaries) - 1)]
                w = iv >> 5
                ok &= w < (hi - lo)
                words = self.cat_threshold[np.where(ok, lo + w, 0)] if len(self.cat_threshold) else np.zeros(len(idx), dtype=np.uint64)
                hit = ok & (((words >> (iv & 31).astype(np.uint64)) & np.uint64(1)) != 0)
                go_left = np.where(is_cat, hit, go_left)
            nxt = np.where(go_left, self.left_child[nd], self.right_child[nd])
            node[idx] = nxt
            active[idx] = nxt >= 0
        return ~node

    def predict(self, X):
        return self.leaf_value[self.predict_leaf(X)]


class TextModel:
    def __init__(self, model_str):
        self.header = {}
        self.average_output = False
        self.trees = []
        self.parameters = {}
        blocks = model_str.split("\n\n")
        for b in blocks:
            lines = [l for l in b.strip().split("\n") if l]
            if not lines:
                continue
            if lines[0].startswith("Tree="):
                kv = dict(l.split("=", 1) for l in lines[1:] if "=" in l)
                self.trees.append(TextTree(kv))
            elif lines[0] == "tree":
                self.header = dict(l.split("=", 1) for l in lines[1:] if "=" in l)
                self.average_output = "average_output" in lines      # boosting=rf: predictions are the mean of the iterations' trees
            elif lines[0] == "parameters:":
                for l in lines[1:]:
                    if l.startswith("[") and ": " in l:
                        k, v = l[1:-1].split(": ", 1)
                        self.parameters[k] = v
        self.objective = self.header.get("objective", "")
        self.num_class = int(self.header.get("num_class", 1))
        self.num_tree_per_iteration = int(self.header.get("num_tree_per_iteration", 1))      # tree t belongs to class t % K
        self.max_feature_idx = int(self.header.get("max_feature_idx", -1))
        self.feature_names = self.header.get("feature_names", "").split()

    def predict_raw(self, X, num_iteration=None):
        """Raw scores: [n] for one tree per iteration, [n, K] for K (multiclass)."""
        X = np.asarray(X, dtype=np.float64)
        K = self.num_tree_per_iteration
        out = np.zeros((X.shape[0], K), dtype=np.float64)
        trees = self.trees if num_iteration is None else self.trees[:num_iteration * K]
        for i, t in enumerate(trees):
            out[:, i % K] += t.predict(X)
        if self.average_output and trees:
            out /= len(trees) // K
        return out[:, 0] if K == 1 else out

    def predict(self, X):
        raw = self.predict_raw(X)
        if self.objective.startswith("multiclassova"):      # MulticlassOVA::ConvertOutput: one sigmoid per class
            sig = 1.0
            for tok in self.objective.split():
                if tok.startswith("sigmoid:"):
                    sig = float(tok.split(":")[1])
            return 1.0 / (1.0 + np.exp(-sig * raw))
        if self.objective.startswith("multiclass"):
            e = np.exp(raw - raw.max(axis=1, keepdims=True))
            return e / e.sum(axis=1, keepdims=True)
        if self.objective.startswith("binary"):
            sig = 1.0
            for tok in self.objective.split():
                if tok.startswith("sigmoid:"):
                    sig = float(tok.split(":")[1])
            return 1.0 / (1.0 + np.exp(-sig * raw))
        return raw

    def structure(self):
        """(split_feature, threshold, left_child, right_child) per tree — what 'identical trees' means in float mode."""
        return [(t.split_feature.tolist(), t.threshold.tolist(), t.left_child.tolist(), t.right_child.tolist()) if t.num_leaves > 1 else None
                for t in self.trees]
