"""Model string loader / writer through the C ABI without a GPU: LGBM_BoosterLoadModelFromString parses
what LGBM_BoosterSaveModelToString writes, and writing it again gives the same text."""
import ctypes as C
import os

import pytest

from helpers import strip_parameters

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def _save(native, h):
    size = 1 << 22
    buf = C.create_string_buffer(size)
    n = C.c_int64()
    assert native.LGBM_BoosterSaveModelToString(h, 0, -1, 0, C.c_int64(size), C.byref(n), buf) == 0
    assert n.value <= size
    return buf.value.decode()


def test_load_then_save_reproduces_the_golden_model(native):
    text = open(os.path.join(GOLDEN, "quant_model.txt")).read()
    h, it = C.c_void_p(), C.c_int()
    assert native.LGBM_BoosterLoadModelFromString(text.encode(), C.byref(it), C.byref(h)) == 0, native.LGBM_GetLastError()
    assert it.value == text.count("Tree=")
    for fn, want in [(native.LGBM_BoosterNumberOfTotalModel, it.value), (native.LGBM_BoosterGetNumFeature, 28), (native.LGBM_BoosterGetNumClasses, 1)]:
        out = C.c_int()
        assert fn(h, C.byref(out)) == 0 and out.value == want
    assert strip_parameters(_save(native, h)) == text
    # feature importances from the loaded trees (split counts are the model text's feature_importances block)
    import numpy as np
    imp = np.zeros(28)
    assert native.LGBM_BoosterFeatureImportance(h, -1, 0, imp.ctypes.data_as(C.POINTER(C.c_double))) == 0
    block = text.split("feature_importances:\n")[1].split("\n\n")[0].strip().splitlines()
    want = {int(l.split("=")[0].split("_")[1]): int(l.split("=")[1]) for l in block}
    assert {j: int(v) for j, v in enumerate(imp) if v > 0} == want
    gain = np.zeros(28)
    assert native.LGBM_BoosterFeatureImportance(h, -1, 1, gain.ctypes.data_as(C.POINTER(C.c_double))) == 0
    assert (gain > 0).sum() == len(want) and gain.sum() > 0
    # a loaded booster cannot train
    fin = C.c_int()
    assert native.LGBM_BoosterUpdateOneIter(h, C.byref(fin)) == -1
    assert b"only predict" in native.LGBM_GetLastError()
    assert native.LGBM_BoosterFree(h) == 0


@pytest.mark.parametrize("bad", ["", "not a model", "tree\nnum_class=3\n", "tree\nmax_feature_idx=1\nTree=0\nnum_leaves=2\nsplit_feature=5\n",
                                 "tree\nmax_feature_idx=1\nTree=0\nnum_leaves=2\nnum_cat=1\n"])
def test_malformed_model_strings_are_rejected(native, bad):
    h, it = C.c_void_p(), C.c_int()
    assert native.LGBM_BoosterLoadModelFromString(bad.encode(), C.byref(it), C.byref(h)) == -1
    assert len(native.LGBM_GetLastError()) > 0


def test_dump_model_json_describes_the_same_trees(native):
    """LGBM_BoosterDumpModel (LightGBMBooster.scala:480-499): valid JSON in upstream's DumpModel layout whose nested nodes are the
    text model's arrays."""
    import json

    import numpy as np

    from synapseml_b200.lightgbm.booster import LightGBMBooster
    from synapseml_b200.lightgbm.model_text import TextModel
    text = open(os.path.join(GOLDEN, "quant_model.txt")).read()
    lb = LightGBMBooster(model_str=text)
    doc = json.loads(lb.dump_model())
    tm = TextModel(text)
    assert doc["name"] == "tree" and doc["num_class"] == 1 and doc["max_feature_idx"] == 27
    assert doc["objective"] == "binary sigmoid:1" and len(doc["feature_names"]) == 28
    assert set(doc["feature_infos"]) == set(doc["feature_names"]) and doc["feature_infos"]["Column_0"]["min_value"] < doc["feature_infos"]["Column_0"]["max_value"]
    assert len(doc["tree_info"]) == len(tm.trees)

    def walk(node, t, seen):
        if "leaf_index" in node:
            i = node["leaf_index"]
            assert node["leaf_value"] == t.leaf_value[i] and node["leaf_count"] == t.leaf_count[i]
            seen["leaves"] += 1
            return ~i
        k = node["split_index"]
        assert node["split_feature"] == t.split_feature[k] and node["threshold"] == t.threshold[k]
        assert node["internal_count"] == t.internal_count[k] and node["decision_type"] == "<="
        assert node["default_left"] == bool(t.decision_type[k] & 2)
        assert node["missing_type"] == ("None", "Zero", "NaN")[(t.decision_type[k] >> 2) & 3]
        assert walk(node["left_child"], t, seen) == t.left_child[k] and walk(node["right_child"], t, seen) == t.right_child[k]
        return k

    for info, t in zip(doc["tree_info"], tm.trees):
        seen = {"leaves": 0}
        assert info["num_leaves"] == t.num_leaves and info["num_cat"] == 0
        walk(info["tree_structure"], t, seen)
        assert seen["leaves"] == t.num_leaves
    block = text.split("feature_importances:\n")[1].split("\n\n")[0].strip().splitlines()
    assert {k: int(v) for k, v in doc["feature_importances"].items()} == {l.split("=")[0]: int(l.split("=")[1]) for l in block}
    lb.free_native_memory()


CAT_MODEL = """tree
version=v3
num_class=1
num_tree_per_iteration=1
label_index=0
max_feature_idx=1
objective=regression
feature_names=Column_0 Column_1
feature_infos=-1:2:0:35 [0:1]
tree_sizes=1

Tree=0
num_leaves=3
num_cat=1
split_feature=0 1
split_gain=10 2
threshold=0 0.5
decision_type=1 2
left_child=1 -1
right_child=-2 -3
leaf_value=1 2 3
leaf_weight=5 5 10
leaf_count=5 5 10
internal_value=0 0
internal_weight=20 10
internal_count=20 10
cat_boundaries=0 2
cat_threshold=5 8
is_linear=0
shrinkage=1


end of trees

feature_importances:
Column_0=1
Column_1=1

parameters:
[objective: regression]
end of parameters

pandas_categorical:null
"""


def test_categorical_model_text_round_trip_and_dump(native):
    """A model with a categorical split (Tree::SplitCategorical layout: num_cat, decision_type bit 0, cat_boundaries, cat_threshold
    bitsets over the category values) is parsed, written back unchanged and dumped as upstream's JSON ("==", categories joined by
    "||") — all on the host."""
    import json
    h, it = C.c_void_p(), C.c_int()
    assert native.LGBM_BoosterLoadModelFromString(CAT_MODEL.encode(), C.byref(it), C.byref(h)) == 0, native.LGBM_GetLastError()
    once = _save(native, h)
    for frag in ("num_cat=1", "decision_type=1 2", "cat_boundaries=0 2", "cat_threshold=5 8", "feature_infos=-1:2:0:35 [0:1]"):
        assert frag in once, frag
    size = 1 << 20
    buf, n = C.create_string_buffer(size), C.c_int64()
    assert native.LGBM_BoosterDumpModel(h, 0, -1, 0, C.c_int64(size), C.byref(n), buf) == 0
    dump = json.loads(buf.value.decode())
    root = dump["tree_info"][0]["tree_structure"]
    # words 5 (bits 0, 2) and 8 (bit 3 -> category 35): the categories that go left
    assert root["decision_type"] == "==" and root["threshold"] == "0||2||35" and dump["tree_info"][0]["num_cat"] == 1
    assert root["left_child"]["decision_type"] == "<=" and root["right_child"]["leaf_index"] == 1
    assert dump["feature_infos"]["Column_0"] == {"min_value": -1, "max_value": 35, "values": [-1, 2, 0, 35]}
    assert native.LGBM_BoosterFree(h) == 0
    h2 = C.c_void_p()
    assert native.LGBM_BoosterLoadModelFromString(once.encode(), C.byref(it), C.byref(h2)) == 0
    assert _save(native, h2) == once
    assert native.LGBM_BoosterFree(h2) == 0
    # a categorical node that points outside the tree's category sets is refused
    bad = CAT_MODEL.replace("threshold=0 0.5", "threshold=3 0.5")
    assert native.LGBM_BoosterLoadModelFromString(bad.encode(), C.byref(it), C.byref(h2)) == -1
