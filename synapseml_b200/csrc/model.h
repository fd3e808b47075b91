// Host-side tree container and the LightGBM text model (format: SURVEY.md A.10).  The text model is the only output of
// training the reference reads back: LGBM_BoosterSaveModelToStringSWIG at
// lightgbm/src/main/scala/com/microsoft/azure/synapse/ml/lightgbm/booster/LightGBMBooster.scala:272-278
// (called from BasePartitionTask.scala:450-461).
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace lgb {

struct HostTree {
  int num_leaves = 1;
  double shrinkage = 1.0;
  std::vector<int> left_child, right_child, split_feature_inner, split_feature, leaf_count, internal_count, leaf_depth;
  std::vector<int> threshold_in_bin;
  std::vector<int8_t> decision_type;
  std::vector<float> split_gain;
  std::vector<double> threshold, leaf_value, leaf_weight, internal_value, internal_weight;
  // categorical splits (decision_type bit 0): node i of that kind keeps its index k among them in threshold[i]; its categories are
  // the bits set in words [cat_boundaries[k], cat_boundaries[k + 1]) of cat_threshold (Tree::SplitCategorical, upstream tree.h).
  // cat_bits_inner: for trees trained here, the same sets over BINS, 8 words per internal node (empty: no categorical split).
  int num_cat = 0;
  std::vector<int> cat_boundaries{0};
  std::vector<uint32_t> cat_threshold;
  std::vector<uint32_t> cat_bits_inner;
  // Tree::CategoricalDecision: NaN, negative and unlisted values go right
  bool CategoricalGoesLeft(int node, double fval) const;

  void Shrinkage(double rate);
  void AddBias(double val);
  void AsConstantTree(double val);
  std::string ToString() const;
  int GetLeaf(const double* row) const;
  double Predict(const double* row) const { return leaf_value[GetLeaf(row)]; }
};

struct ModelHeader {
  int max_feature_idx = 0;
  bool average_output = false;                          // boosting=rf: a prediction is the mean of the iterations' trees
  int num_class = 1, num_tree_per_iteration = 1;      // multiclass: one tree per class and iteration, class k's trees are k, k + K, ...
  std::string objective;
  std::vector<std::string> feature_names;
  std::vector<std::string> feature_infos;
  std::vector<std::pair<std::string, std::string>> parameters;
};

// start_iteration / num_iteration count ITERATIONS (num_tree_per_iteration trees each)
std::string ModelToString(const ModelHeader& h, const std::vector<HostTree>& trees, int start_iteration, int num_iteration);
// JSON dump of the same model (upstream GBDT::DumpModel / Tree::ToJSON): what LGBM_BoosterDumpModel returns to
// lightgbm/src/main/scala/com/microsoft/azure/synapse/ml/lightgbm/booster/LightGBMBooster.scala:480-499 (dumpModel).
// importance: per real feature, already computed by the caller (split counts or total gains).
std::string ModelToJson(const ModelHeader& h, const std::vector<HostTree>& trees, int start_iteration, int num_iteration,
                        const std::vector<double>& importance);

// Inverse of ModelToString for the subset this library writes (no linear trees):
// what LGBM_BoosterLoadModelFromString reads when the Scala side rebuilds a booster from the model string on an executor
// (lightgbm/src/main/scala/com/microsoft/azure/synapse/ml/lightgbm/booster/LightGBMBooster.scala:41-49).  Throws
// std::runtime_error on malformed input.
void ParseModelString(const std::string& text, ModelHeader* h, std::vector<HostTree>* trees);

}  // namespace lgb
