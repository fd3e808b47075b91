// Parameter string ("k=v k=v ...") -> Config.  This is the native end of the strings the reference builds at
// lightgbm/src/main/scala/com/microsoft/azure/synapse/ml/lightgbm/params/BaseTrainParams.scala:25-40 (booster) and
// lightgbm/.../LightGBMBase.scala:374-386 (dataset).  First occurrence of a key wins, which is what makes the reference's
// passThroughArgs override explicit setters (core/.../utils/ParamsStringBuilder.scala:100-109).
#pragma once
#include <map>
#include <string>
#include <vector>

namespace lgb {

// Upstream's include/LightGBM/utils/random.h: a 32-bit LCG (x = 214013 x + 2531011), 15-bit and 31-bit draws.  Host only.
struct HostRandom {
  unsigned int x;
  explicit HostRandom(int seed) : x(static_cast<unsigned int>(seed)) {}
  int RandInt16() { x = 214013u * x + 2531011u; return static_cast<int>((x >> 16) & 0x7FFF); }
  int RandInt32() { x = 214013u * x + 2531011u; return static_cast<int>(x & 0x7FFFFFFF); }
  int NextShort(int lo, int hi) { return RandInt16() % (hi - lo) + lo; }
  int NextInt(int lo, int hi) { return RandInt32() % (hi - lo) + lo; }
  float NextFloat() { return static_cast<float>(RandInt16()) / 32768.0f; }
  // K of 0..N-1, ascending (Random::Sample): sequential selection sampling for large K, Floyd's set sampling for small K
  std::vector<int> Sample(int N, int K);
};

struct Config {
  std::string objective = "regression";
  std::string boosting = "gbdt";
  std::string tree_learner = "serial";
  int num_iterations = 100;
  double learning_rate = 0.1;
  int num_leaves = 31;
  int max_depth = -1;
  int max_bin = 255;
  int min_data_in_bin = 3;
  int bin_construct_sample_cnt = 200000;
  int min_data_in_leaf = 20;
  double min_sum_hessian_in_leaf = 1e-3;
  double lambda_l1 = 0.0, lambda_l2 = 0.0, min_gain_to_split = 0.0, max_delta_step = 0.0;
  bool boost_from_average = true;
  double sigmoid = 1.0;
  bool is_unbalance = false;
  double scale_pos_weight = 1.0;
  bool use_missing = true, zero_as_missing = false, feature_pre_filter = true;
  bool use_quantized_grad = false;
  int num_grad_quant_bins = 4;
  int num_class = 1;
  bool stochastic_rounding = true;
  bool quant_train_renew_leaf = false;
  int seed = 0;
  // column / row sampling (upstream: src/treelearner/col_sampler.hpp, src/boosting/bagging.hpp).  `seed` derives the specific
  // seeds in upstream's Config::Set order unless they are given themselves.
  double alpha = 0.9, fair_c = 1.0, poisson_max_delta_step = 0.7, tweedie_variance_power = 1.5;      // huber / fair / poisson / tweedie
  double feature_fraction = 1.0, feature_fraction_bynode = 1.0;
  int feature_fraction_seed = 2;
  double bagging_fraction = 1.0, pos_bagging_fraction = 1.0, neg_bagging_fraction = 1.0;
  int bagging_freq = 0, bagging_seed = 3;
  // boosting = dart / goss (upstream config.h defaults)
  double drop_rate = 0.1, skip_drop = 0.5, top_rate = 0.2, other_rate = 0.1;
  int max_drop = 50, drop_seed = 4;
  int objective_seed = 5;      // rank_xendcg's per-query streams (LightGBMParams.scala objectiveSeed)
  bool uniform_drop = false, xgboost_dart_mode = false;
  int num_machines = 1;
  int num_threads = 0;
  int verbosity = 1;
  int lambdarank_truncation_level = 30;
  bool lambdarank_norm = true;
  std::vector<double> label_gain;
  std::vector<int> eval_at;
  // categorical features: column indexes (LightGBMBase.scala:162-198 resolves slot names to indexes before the native call) and the
  // knobs of the categorical split search (upstream config.h defaults; CategoricalParams in params/TrainParams.scala)
  std::vector<int> categorical_feature;
  std::vector<int> max_bin_by_feature;      // one max_bin per column (dataset construction only, as upstream)
  std::vector<int> monotone_constraints;    // per column: +1 increasing, -1 decreasing, 0 free (LightGBMParams.scala monotoneConstraints)
  std::string monotone_constraints_method = "basic";
  double monotone_penalty = 0.0;
  int max_cat_to_onehot = 4, max_cat_threshold = 32, min_data_per_group = 100;
  double cat_smooth = 10.0, cat_l2 = 10.0;
  std::map<std::string, std::string> raw;   // canonical key -> value as received (echoed into the model's parameters block)

  void Set(const std::string& params);      // merges; keys already present in `params` earlier win over later ones
  static std::string CanonicalKey(const std::string& k);
};

}  // namespace lgb
