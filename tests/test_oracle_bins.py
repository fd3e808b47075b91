"""Oracle vs hand-derived known answers for bin finding.  The derivations are written out so the
expected values do not come from the code under test."""
import json
import os

import numpy as np

KZ = float(np.float32(1e-35))
INF = float("inf")
up = lambda x: float(np.nextafter(x, np.inf))
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def test_few_distinct_with_implicit_zeros(oracle):
    # 10 sampled rows: 4 zeros (absent from the sample arrays), 1.0 x3, 2.0 x3.  distinct = [0,1,2], counts = [4,3,3].
    # No negative part; zero gets its own bin closed at +1e-35; right part has 2 distinct <= max_bin:
    # cur=3 >= min_data_in_bin=3 -> bound nextafter(1.5); last bound +inf.
    r = oracle.find_bin([1.0, 2.0, 1.0, 2.0, 1.0, 2.0], total_sample_cnt=10)
    assert r["bounds"].tolist() == [KZ, up(1.5), INF]
    assert (r["num_bin"], r["missing_type"], r["default_bin"], r["most_freq_bin"], r["is_trivial"]) == (3, 0, 0, 0, False)


def test_negative_and_positive_zero_crossing(oracle):
    # -2 x5, 3 x5, no zeros: a zero entry with count 0 is still inserted at the sign change -> distinct [-2,0,3], counts [5,0,5].
    # left part -> single bound replaced by -1e-35; zero bin closed at +1e-35; right part -> +inf.
    # cnt_in_bin = [5,0,5]; default_bin = bin of 0 = 1; argmax = 0 but its share 0.5 < 0.7 -> most_freq_bin = default_bin.
    r = oracle.find_bin([-2.0] * 5 + [3.0] * 5, total_sample_cnt=10)
    assert r["bounds"].tolist() == [-KZ, KZ, INF]
    assert (r["default_bin"], r["most_freq_bin"]) == (1, 1)


def test_nan_gets_last_bin(oracle):
    # 1 x3, 2 x3, NaN x2 (8 rows, no zeros): missing_type NaN; bins found with max_bin-1 on the 6 non-NaN rows, then a NaN bound.
    r = oracle.find_bin([1.0, 1.0, 1.0, 2.0, 2.0, 2.0, float("nan"), float("nan")], total_sample_cnt=8)
    b = r["bounds"]
    assert b[:3].tolist() == [KZ, up(1.5), INF] and np.isnan(b[3])
    # cnt_in_bin = [0,3,3,2]: argmax = 1 != default_bin 0 and 3/8 < 0.7 -> most_freq_bin = 0
    assert (r["num_bin"], r["missing_type"], r["default_bin"], r["most_freq_bin"]) == (4, 2, 0, 0)


def test_greedy_more_distinct_than_bins(oracle):
    # 1..20 once each, max_bin=4 -> 3 bins for the positive part, mean 20/3: bins close after 7 values (cur=7 >= 6.67) and after
    # the next 7 (mean becomes 13/2 = 6.5) -> bounds between 7|8 and 14|15.
    r = oracle.find_bin(np.arange(1, 21, dtype=np.float64), total_sample_cnt=20, max_bin=4, min_data_in_bin=1)
    assert r["bounds"].tolist() == [KZ, up(7.5), up(14.5), INF]


def test_big_count_value_is_pinned(oracle):
    # values 1..10 once each and 5.0 forty more times (50 rows), max_bin=4 -> 3 bins, mean0 = 16.67 -> 5.0 (41) is "big":
    # rest 9 samples over 2 bins (mean 4.5).  Sweep: i=3 (value 4): next is big and cur=4 >= max(1, 2.25) -> close at 4|5;
    # i=4 (value 5) big -> close at 5|6; then bin_cnt reaches max_bin-1.
    vals = list(range(1, 11)) + [5.0] * 40
    r = oracle.find_bin(np.array(vals, dtype=np.float64), total_sample_cnt=50, max_bin=4, min_data_in_bin=1)
    assert r["bounds"].tolist() == [KZ, up(4.5), up(5.5), INF]


def test_trivial_feature(oracle):
    # one distinct value and no zeros: bounds [+1e-35, inf] (an empty zero bin + one value bin), cnt_in_bin = [0, 10].
    # No cut has >= 1 sample on both sides, so the pre-filter drops it as soon as the filter count is >= 1; with a filter
    # count of 0 (min_data_in_leaf * sample / N rounds to 0, as in the 10M-row config) it is kept with 2 bins.
    r = oracle.find_bin([4.25] * 10, total_sample_cnt=10, min_split_data=1)
    assert r["is_trivial"] and r["num_bin"] == 2
    r = oracle.find_bin([4.25] * 10, total_sample_cnt=10, min_split_data=0)
    assert not r["is_trivial"] and r["bounds"].tolist() == [KZ, INF]
    # no non-zero value at all: a single zero entry -> one bin -> trivial regardless of the filter
    r = oracle.find_bin([], total_sample_cnt=10, min_split_data=0)
    assert r["is_trivial"] and r["num_bin"] == 1


def test_pre_filter_drops_unsplittable_feature(oracle):
    # 2 non-zero values out of 1000 rows with filter count 5: no cut leaves >= 5 rows on both sides -> trivial
    r = oracle.find_bin([1.0, 1.0], total_sample_cnt=1000, min_split_data=5, min_data_in_bin=1)
    assert r["is_trivial"]
    r = oracle.find_bin([1.0] * 10, total_sample_cnt=1000, min_split_data=5, min_data_in_bin=1)
    assert not r["is_trivial"]


def test_value_to_bin_is_first_bin_whose_upper_bound_covers_the_value(oracle):
    X = np.array([[1.0], [1.0], [1.0], [2.0], [2.0], [2.0], [np.nan], [np.nan], [1.5], [1.6], [100.0], [0.0], [-5.0]])
    ref = oracle.Dataset.from_sample(X[:8], 1, 8, "max_bin=255 min_data_in_leaf=1")
    ds = ref.create_by_reference(len(X))
    ds.init_streaming()
    ds.push_rows(X, 0, np.zeros(len(X), dtype=np.float32))
    ds.mark_finished()
    assert ds.bin_column(0).tolist() == [1, 1, 1, 2, 2, 2, 3, 3, 1, 2, 2, 0, 0]


def test_golden_bin_mappers(oracle):
    """The committed fixture (tests/golden/make_golden.py) pins the oracle against accidental change; the GPU tests compare the
    CUDA K1/K2 kernels with the same file."""
    from helpers import edge_case_matrix
    with open(os.path.join(GOLDEN, "edge_bins.json")) as f:
        g = json.load(f)
    X = edge_case_matrix(g["n"], g["seed"])
    ref = oracle.Dataset.from_sample(X[:g["sample"]], 1, g["n"], g["params"])
    for j, feat in enumerate(g["features"]):
        info = ref.feature_info(j)
        assert info["num_bin"] == feat["num_bin"] and info["missing_type"] == feat["missing_type"]
        assert info["default_bin"] == feat["default_bin"] and info["most_freq_bin"] == feat["most_freq_bin"]
        assert info["is_trivial"] == feat["is_trivial"]
        got = ref.bin_upper_bounds(j)
        exp = np.array([float(x) for x in feat["bounds_hex_as_float"]])
        assert [float.hex(float(x)) if not np.isnan(x) else "nan" for x in got] == feat["bounds_hex"]
        del exp
