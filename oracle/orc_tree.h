// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_common.h header; parity unpinned).
//
// Tree container and LightGBM text-model block.  Restates upstream Tree::Split / SplitCategorical / Shrinkage / AddBias / ToString
// (src/io/tree.cpp, include/LightGBM/tree.h) and Tree::NumericalDecision / CategoricalDecision, SURVEY.md A.8/A.10.
// Reference consumer: LGBM_BoosterSaveModelToStringSWIG at lightgbm/.../booster/LightGBMBooster.scala:272-278.
#pragma once
#include <cstdio>

#include "orc_common.h"

namespace orc {

inline std::string FmtG17(double v) { char b[64]; std::snprintf(b, sizeof(b), "%.17g", v); return b; }
inline std::string FmtG(double v) { char b[64]; std::snprintf(b, sizeof(b), "%g", v); return b; }

inline double MaybeRoundToZero(double v) { return (v >= -kZeroThreshold && v <= kZeroThreshold) ? 0.0 : v; }

struct Tree {
  int max_leaves;
  int num_leaves = 1;
  std::vector<int> left_child, right_child, split_feature_inner, split_feature, leaf_parent, leaf_depth;
  std::vector<uint32_t> threshold_in_bin;
  std::vector<double> threshold;
  std::vector<int8_t> decision_type;
  std::vector<float> split_gain;
  std::vector<double> leaf_value, leaf_weight, internal_value, internal_weight;
  std::vector<int> leaf_count, internal_count;
  double shrinkage = 1.0;
  // categorical splits: node i of that kind stores its index among them in threshold[i]; its categories are the bits set in words
  // [cat_boundaries[k], cat_boundaries[k + 1]) of cat_threshold (category values) / ..._inner (bins)
  int num_cat = 0;
  std::vector<int> cat_boundaries{0}, cat_boundaries_inner{0};
  std::vector<uint32_t> cat_threshold, cat_threshold_inner;

  static bool FindInBitset(const uint32_t* bits, int n, int pos) {
    const int i1 = pos / 32;
    if (i1 >= n) return false;
    return (bits[i1] >> (pos % 32)) & 1;
  }
  static std::vector<uint32_t> ConstructBitset(const int* vals, int n) {
    std::vector<uint32_t> ret;
    for (int i = 0; i < n; ++i) {
      const int i1 = vals[i] / 32, i2 = vals[i] % 32;
      if (static_cast<int>(ret.size()) < i1 + 1) ret.resize(i1 + 1, 0);
      ret[i1] |= (1u << i2);
    }
    return ret;
  }

  explicit Tree(int max_leaves_) : max_leaves(max_leaves_) {
    int n = std::max(max_leaves, 2);
    left_child.assign(n - 1, 0); right_child.assign(n - 1, 0);
    split_feature_inner.assign(n - 1, 0); split_feature.assign(n - 1, 0);
    threshold_in_bin.assign(n - 1, 0); threshold.assign(n - 1, 0.0);
    decision_type.assign(n - 1, 0); split_gain.assign(n - 1, 0.0f);
    internal_value.assign(n - 1, 0.0); internal_weight.assign(n - 1, 0.0); internal_count.assign(n - 1, 0);
    leaf_parent.assign(n, -1); leaf_depth.assign(n, 0);
    leaf_value.assign(n, 0.0); leaf_weight.assign(n, 0.0); leaf_count.assign(n, 0);
  }

  int Split(int leaf, int feature_inner, int real_feature, uint32_t threshold_bin, double threshold_double,
            double left_value, double right_value, int left_cnt, int right_cnt, double left_weight, double right_weight,
            float gain, int missing_type, bool default_left) {
    int new_node = num_leaves - 1;
    int parent = leaf_parent[leaf];
    if (parent >= 0) {
      if (left_child[parent] == ~leaf) left_child[parent] = new_node; else right_child[parent] = new_node;
    }
    split_feature_inner[new_node] = feature_inner;
    split_feature[new_node] = real_feature;
    split_gain[new_node] = gain;
    left_child[new_node] = ~leaf;
    right_child[new_node] = ~num_leaves;
    leaf_parent[leaf] = new_node;
    leaf_parent[num_leaves] = new_node;
    internal_weight[new_node] = left_weight + right_weight;
    internal_value[new_node] = leaf_value[leaf];
    internal_count[new_node] = left_cnt + right_cnt;
    leaf_value[leaf] = std::isnan(left_value) ? 0.0 : left_value;
    leaf_weight[leaf] = left_weight;
    leaf_count[leaf] = left_cnt;
    leaf_value[num_leaves] = std::isnan(right_value) ? 0.0 : right_value;
    leaf_weight[num_leaves] = right_weight;
    leaf_count[num_leaves] = right_cnt;
    leaf_depth[num_leaves] = leaf_depth[leaf] + 1;
    leaf_depth[leaf]++;
    int8_t dt = 0;                       // bit0 categorical = 0
    if (default_left) dt |= 2;           // bit1 default_left
    dt |= static_cast<int8_t>((missing_type & 3) << 2);
    decision_type[new_node] = dt;
    threshold_in_bin[new_node] = threshold_bin;
    threshold[new_node] = threshold_double;
    ++num_leaves;
    return num_leaves - 1;
  }

  // Tree::SplitCategorical: bitsets over bins (threshold_bin) and over category values (threshold_cat); default_left is false
  int SplitCategorical(int leaf, int feature_inner, int real_feature, const std::vector<uint32_t>& threshold_bin,
                       const std::vector<uint32_t>& threshold_cat, double left_value, double right_value, int left_cnt, int right_cnt,
                       double left_weight, double right_weight, float gain, int missing_type) {
    const int new_node = num_leaves - 1;
    Split(leaf, feature_inner, real_feature, 0, 0.0, left_value, right_value, left_cnt, right_cnt, left_weight, right_weight, gain,
          missing_type, false);
    decision_type[new_node] = static_cast<int8_t>(1 | ((missing_type & 3) << 2));
    threshold_in_bin[new_node] = static_cast<uint32_t>(num_cat);
    threshold[new_node] = num_cat;
    ++num_cat;
    cat_boundaries.push_back(cat_boundaries.back() + static_cast<int>(threshold_cat.size()));
    cat_threshold.insert(cat_threshold.end(), threshold_cat.begin(), threshold_cat.end());
    cat_boundaries_inner.push_back(cat_boundaries_inner.back() + static_cast<int>(threshold_bin.size()));
    cat_threshold_inner.insert(cat_threshold_inner.end(), threshold_bin.begin(), threshold_bin.end());
    return num_leaves - 1;
  }

  void Shrinkage(double rate) {
    for (int i = 0; i < num_leaves - 1; ++i) {
      leaf_value[i] = MaybeRoundToZero(leaf_value[i] * rate);
      internal_value[i] = MaybeRoundToZero(internal_value[i] * rate);
    }
    leaf_value[num_leaves - 1] = MaybeRoundToZero(leaf_value[num_leaves - 1] * rate);
    shrinkage *= rate;
  }

  void AddBias(double val) {
    for (int i = 0; i < num_leaves - 1; ++i) {
      leaf_value[i] = MaybeRoundToZero(leaf_value[i] + val);
      internal_value[i] = MaybeRoundToZero(internal_value[i] + val);
    }
    leaf_value[num_leaves - 1] = MaybeRoundToZero(leaf_value[num_leaves - 1] + val);
    shrinkage = 1.0;
  }

  void AsConstantTree(double val) { num_leaves = 1; shrinkage = 1.0; leaf_value[0] = val; }

  int GetLeaf(const double* row) const {
    if (num_leaves <= 1) return 0;
    int node = 0;
    while (node >= 0) {
      double fval = row[split_feature[node]];
      int mt = (decision_type[node] >> 2) & 3;
      bool dl = (decision_type[node] & 2) != 0;
      if (decision_type[node] & 1) {      // CategoricalDecision: NaN, negative and unlisted values go right
        bool left = false;
        if (!std::isnan(fval) && fval < 2147483648.0) {
          const int iv = static_cast<int>(fval);
          const int k = static_cast<int>(threshold[node]);
          left = iv >= 0 && FindInBitset(cat_threshold.data() + cat_boundaries[k], cat_boundaries[k + 1] - cat_boundaries[k], iv);
        }
        node = left ? left_child[node] : right_child[node];
        continue;
      }
      if (std::isnan(fval) && mt != MISSING_NAN) fval = 0.0;
      if ((mt == MISSING_ZERO && fval >= -kZeroThreshold && fval <= kZeroThreshold) || (mt == MISSING_NAN && std::isnan(fval))) {
        node = dl ? left_child[node] : right_child[node];
      } else {
        node = (fval <= threshold[node]) ? left_child[node] : right_child[node];
      }
    }
    return ~node;
  }
  double Predict(const double* row) const { return leaf_value[GetLeaf(row)]; }

  template <typename T, typename F>
  static std::string Join(const std::vector<T>& v, int n, F fmt) {
    std::string s;
    for (int i = 0; i < n; ++i) { if (i) s += ' '; s += fmt(v[i]); }
    return s;
  }

  std::string ToString() const {
    auto fi = [](int v) { return std::to_string(v); };
    auto fu = [](uint32_t v) { return std::to_string(v); };
    auto f17 = [](double v) { return FmtG17(v); };
    auto fg = [](float v) { return FmtG(v); };
    auto fd = [](int8_t v) { return std::to_string(static_cast<int>(v)); };
    std::string s;
    s += "num_leaves=" + std::to_string(num_leaves) + "\n";
    s += "num_cat=" + std::to_string(num_cat) + "\n";
    if (num_leaves > 1) {
      int ni = num_leaves - 1;
      s += "split_feature=" + Join(split_feature, ni, fi) + "\n";
      s += "split_gain=" + Join(split_gain, ni, fg) + "\n";
      s += "threshold=" + Join(threshold, ni, f17) + "\n";
      s += "decision_type=" + Join(decision_type, ni, fd) + "\n";
      s += "left_child=" + Join(left_child, ni, fi) + "\n";
      s += "right_child=" + Join(right_child, ni, fi) + "\n";
      s += "leaf_value=" + Join(leaf_value, num_leaves, f17) + "\n";
      s += "leaf_weight=" + Join(leaf_weight, num_leaves, f17) + "\n";
      s += "leaf_count=" + Join(leaf_count, num_leaves, fi) + "\n";
      s += "internal_value=" + Join(internal_value, ni, f17) + "\n";
      s += "internal_weight=" + Join(internal_weight, ni, f17) + "\n";
      s += "internal_count=" + Join(internal_count, ni, fi) + "\n";
      if (num_cat > 0) {
        s += "cat_boundaries=" + Join(cat_boundaries, num_cat + 1, fi) + "\n";
        s += "cat_threshold=" + Join(cat_threshold, static_cast<int>(cat_threshold.size()), fu) + "\n";
      }
    } else {
      s += "leaf_value=" + FmtG17(leaf_value[0]) + "\n";
    }
    s += "is_linear=0\n";
    s += "shrinkage=" + FmtG(shrinkage) + "\n";
    s += "\n";
    return s;
  }
};

}  // namespace orc
