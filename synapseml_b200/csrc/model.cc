#include "model.h"

#include <algorithm>
#include <cmath>
#include <cstdio>

namespace lgb {

namespace {
const double kZeroThresholdD = 1e-35f;
inline double RoundTinyToZero(double v) { return (v >= -kZeroThresholdD && v <= kZeroThresholdD) ? 0.0 : v; }
std::string G17(double v) { char b[64]; snprintf(b, sizeof(b), "%.17g", v); return b; }
std::string Gs(double v) { char b[64]; snprintf(b, sizeof(b), "%g", v); return b; }

template <typename T, typename F>
std::string JoinN(const std::vector<T>& v, int n, F f) {
  std::string s;
  for (int i = 0; i < n; ++i) {
    if (i) s.push_back(' ');
    s += f(v[i]);
  }
  return s;
}
}  // namespace

void HostTree::Shrinkage(double rate) {
  for (int i = 0; i < num_leaves - 1; ++i) {
    leaf_value[i] = RoundTinyToZero(leaf_value[i] * rate);
    internal_value[i] = RoundTinyToZero(internal_value[i] * rate);
  }
  leaf_value[num_leaves - 1] = RoundTinyToZero(leaf_value[num_leaves - 1] * rate);
  shrinkage *= rate;
}

void HostTree::AddBias(double val) {
  for (int i = 0; i < num_leaves - 1; ++i) {
    leaf_value[i] = RoundTinyToZero(leaf_value[i] + val);
    internal_value[i] = RoundTinyToZero(internal_value[i] + val);
  }
  leaf_value[num_leaves - 1] = RoundTinyToZero(leaf_value[num_leaves - 1] + val);
  shrinkage = 1.0;
}

void HostTree::AsConstantTree(double val) {
  num_leaves = 1;
  shrinkage = 1.0;
  leaf_value.assign(1, val);
}

int HostTree::GetLeaf(const double* row) const {
  if (num_leaves <= 1) return 0;
  int node = 0;
  while (node >= 0) {
    double fval = row[split_feature[node]];
    const int mt = (decision_type[node] >> 2) & 3;
    const bool default_left = (decision_type[node] & 2) != 0;
    if (std::isnan(fval) && mt != 2) fval = 0.0;
    if ((mt == 1 && fval >= -kZeroThresholdD && fval <= kZeroThresholdD) || (mt == 2 && std::isnan(fval)))
      node = default_left ? left_child[node] : right_child[node];
    else
      node = fval <= threshold[node] ? left_child[node] : right_child[node];
  }
  return ~node;
}

std::string HostTree::ToString() const {
  auto fi = [](int v) { return std::to_string(v); };
  auto fd = [](int8_t v) { return std::to_string(static_cast<int>(v)); };
  auto f17 = [](double v) { return G17(v); };
  auto fg = [](float v) { return Gs(v); };
  std::string s = "num_leaves=" + std::to_string(num_leaves) + "\nnum_cat=0\n";
  if (num_leaves > 1) {
    const int ni = num_leaves - 1;
    s += "split_feature=" + JoinN(split_feature, ni, fi) + "\n";
    s += "split_gain=" + JoinN(split_gain, ni, fg) + "\n";
    s += "threshold=" + JoinN(threshold, ni, f17) + "\n";
    s += "decision_type=" + JoinN(decision_type, ni, fd) + "\n";
    s += "left_child=" + JoinN(left_child, ni, fi) + "\n";
    s += "right_child=" + JoinN(right_child, ni, fi) + "\n";
    s += "leaf_value=" + JoinN(leaf_value, num_leaves, f17) + "\n";
    s += "leaf_weight=" + JoinN(leaf_weight, num_leaves, f17) + "\n";
    s += "leaf_count=" + JoinN(leaf_count, num_leaves, fi) + "\n";
    s += "internal_value=" + JoinN(internal_value, ni, f17) + "\n";
    s += "internal_weight=" + JoinN(internal_weight, ni, f17) + "\n";
    s += "internal_count=" + JoinN(internal_count, ni, fi) + "\n";
  } else {
    s += "leaf_value=" + G17(leaf_value[0]) + "\n";
  }
  s += "is_linear=0\nshrinkage=" + Gs(shrinkage) + "\n\n";
  return s;
}

std::string ModelToString(const ModelHeader& h, const std::vector<HostTree>& trees, int start_iteration, int num_iteration) {
  int first = std::max(0, start_iteration);
  int last = static_cast<int>(trees.size());
  if (num_iteration > 0) last = std::min(last, first + num_iteration);
  std::string s = "tree\nversion=v3\nnum_class=1\nnum_tree_per_iteration=1\nlabel_index=0\n";
  s += "max_feature_idx=" + std::to_string(h.max_feature_idx) + "\n";
  s += "objective=" + h.objective + "\n";
  s += "feature_names=";
  for (size_t j = 0; j < h.feature_names.size(); ++j) { if (j) s.push_back(' '); s += h.feature_names[j]; }
  s += "\nfeature_infos=";
  for (size_t j = 0; j < h.feature_infos.size(); ++j) { if (j) s.push_back(' '); s += h.feature_infos[j]; }
  s += "\n";
  std::vector<std::string> blocks;
  for (int i = first; i < last; ++i) blocks.push_back("Tree=" + std::to_string(i - first) + "\n" + trees[i].ToString() + "\n");
  s += "tree_sizes=";
  for (size_t i = 0; i < blocks.size(); ++i) { if (i) s.push_back(' '); s += std::to_string(blocks[i].size()); }
  s += "\n\n";
  for (auto& b : blocks) s += b;
  s += "end of trees\n\n";
  std::vector<int> imp(h.feature_names.size(), 0);
  for (int i = first; i < last; ++i)
    for (int k = 0; k < trees[i].num_leaves - 1; ++k) imp[trees[i].split_feature[k]]++;
  std::vector<int> order;
  for (size_t j = 0; j < imp.size(); ++j) if (imp[j] > 0) order.push_back(static_cast<int>(j));
  std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return imp[a] > imp[b]; });
  s += "feature_importances:\n";
  for (int j : order) s += h.feature_names[j] + "=" + std::to_string(imp[j]) + "\n";
  s += "\nparameters:\n";
  for (auto& kv : h.parameters) s += "[" + kv.first + ": " + kv.second + "]\n";
  s += "end of parameters\n\npandas_categorical:null\n";
  return s;
}

std::string FormatG17(double v) { return G17(v); }
std::string FormatG(double v) { return Gs(v); }

}  // namespace lgb
