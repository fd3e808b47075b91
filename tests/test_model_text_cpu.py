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
