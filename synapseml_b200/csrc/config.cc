#include "config.h"

#include <algorithm>
#include <cctype>
#include <cmath>
#include <set>
#include <sstream>
#include <stdexcept>

namespace lgb {

std::vector<int> HostRandom::Sample(int N, int K) {
  std::vector<int> ret;
  if (K > N || K <= 0) return ret;
  ret.reserve(K);
  if (K == N) {
    for (int i = 0; i < N; ++i) ret.push_back(i);
  } else if (K > 1 && K > (N / std::log2(static_cast<double>(K)))) {
    for (int i = 0; i < N; ++i) {
      const double prob = (K - static_cast<int>(ret.size())) / static_cast<double>(N - i);
      if (NextFloat() < prob) ret.push_back(i);
    }
  } else {
    std::set<int> chosen;
    for (int r = N - K; r < N; ++r) {
      const int v = NextInt(0, r);
      if (!chosen.insert(v).second) chosen.insert(r);
    }
    ret.assign(chosen.begin(), chosen.end());
  }
  return ret;
}

std::string Config::CanonicalKey(const std::string& k) {
  // alias -> canonical name (subset of upstream's config_auto.cpp alias table that SynapseML users can reach)
  static const char* const kAlias[][2] = {
      {"objective_type", "objective"}, {"app", "objective"}, {"application", "objective"}, {"loss", "objective"},
      {"boosting_type", "boosting"}, {"boost", "boosting"},
      {"tree", "tree_learner"}, {"tree_type", "tree_learner"}, {"tree_learner_type", "tree_learner"},
      {"num_iteration", "num_iterations"}, {"n_iter", "num_iterations"}, {"num_tree", "num_iterations"},
      {"num_trees", "num_iterations"}, {"num_round", "num_iterations"}, {"num_rounds", "num_iterations"},
      {"nrounds", "num_iterations"}, {"num_boost_round", "num_iterations"}, {"n_estimators", "num_iterations"},
      {"max_iter", "num_iterations"},
      {"shrinkage_rate", "learning_rate"}, {"eta", "learning_rate"},
      {"num_leaf", "num_leaves"}, {"max_leaves", "num_leaves"}, {"max_leaf", "num_leaves"}, {"max_leaf_nodes", "num_leaves"},
      {"min_data_per_leaf", "min_data_in_leaf"}, {"min_data", "min_data_in_leaf"}, {"min_child_samples", "min_data_in_leaf"},
      {"min_samples_leaf", "min_data_in_leaf"},
      {"min_sum_hessian_per_leaf", "min_sum_hessian_in_leaf"}, {"min_sum_hessian", "min_sum_hessian_in_leaf"},
      {"min_hessian", "min_sum_hessian_in_leaf"}, {"min_child_weight", "min_sum_hessian_in_leaf"},
      {"reg_alpha", "lambda_l1"}, {"l1_regularization", "lambda_l1"},
      {"reg_lambda", "lambda_l2"}, {"lambda", "lambda_l2"}, {"l2_regularization", "lambda_l2"},
      {"min_split_gain", "min_gain_to_split"},
      {"max_bins", "max_bin"}, {"subsample_for_bin", "bin_construct_sample_cnt"},
      {"unbalance", "is_unbalance"}, {"unbalanced_sets", "is_unbalance"},
      {"num_thread", "num_threads"}, {"nthread", "num_threads"}, {"nthreads", "num_threads"}, {"n_jobs", "num_threads"},
      {"random_seed", "seed"}, {"random_state", "seed"},
      {"num_machine", "num_machines"}, {"verbose", "verbosity"},
      {"ndcg_eval_at", "eval_at"}, {"ndcg_at", "eval_at"},
      {"sub_feature", "feature_fraction"}, {"colsample_bytree", "feature_fraction"},
      {"sub_feature_bynode", "feature_fraction_bynode"}, {"colsample_bynode", "feature_fraction_bynode"},
      {"sub_row", "bagging_fraction"}, {"subsample", "bagging_fraction"}, {"bagging", "bagging_fraction"},
      {"subsample_freq", "bagging_freq"}, {"bagging_fraction_seed", "bagging_seed"},
      {"mc", "monotone_constraints"}, {"monotone_constraint", "monotone_constraints"}, {"monotonic_cst", "monotone_constraints"},
      {"monotone_constraining_method", "monotone_constraints_method"}, {"mc_method", "monotone_constraints_method"},
      {"monotone_splits_penalty", "monotone_penalty"}, {"ms_penalty", "monotone_penalty"}, {"mc_penalty", "monotone_penalty"},
      {"cat_feature", "categorical_feature"}, {"categorical_column", "categorical_feature"},
      {"cat_column", "categorical_feature"}, {"categorical_features", "categorical_feature"},
  };
  for (auto& a : kAlias)
    if (k == a[0]) return a[1];
  return k;
}

static bool ToBool(const std::string& key, std::string v) {
  std::transform(v.begin(), v.end(), v.begin(), [](unsigned char c) { return std::tolower(c); });
  if (v == "true" || v == "1" || v == "+") return true;
  if (v == "false" || v == "0" || v == "-") return false;
  throw std::runtime_error("Parameter " + key + " should be of type bool, got \"" + v + "\"");
}

static double ToDouble(const std::string& key, const std::string& v) {
  try {
    size_t pos = 0;
    double d = std::stod(v, &pos);
    if (pos != v.size()) throw std::invalid_argument(v);
    return d;
  } catch (...) {
    throw std::runtime_error("Parameter " + key + " should be numeric, got \"" + v + "\"");
  }
}

template <typename T>
static std::vector<T> ToList(const std::string& key, const std::string& v) {
  std::vector<T> out;
  std::stringstream ss(v);
  std::string tok;
  while (std::getline(ss, tok, ','))
    if (!tok.empty()) out.push_back(static_cast<T>(ToDouble(key, tok)));
  return out;
}

void Config::Set(const std::string& params) {
  std::map<std::string, std::string> kv;
  std::stringstream ss(params);
  std::string tok;
  while (ss >> tok) {
    size_t eq = tok.find('=');
    if (eq == std::string::npos || eq == 0) continue;
    std::string k = CanonicalKey(tok.substr(0, eq));
    if (!kv.count(k)) kv[k] = tok.substr(eq + 1);
  }
  for (auto& p : kv) raw[p.first] = p.second;
  auto has = [&](const char* k) { return kv.count(k) > 0; };
#define GET_I(name) if (has(#name)) name = static_cast<int>(ToDouble(#name, kv[#name]))
#define GET_D(name) if (has(#name)) name = ToDouble(#name, kv[#name])
#define GET_B(name) if (has(#name)) name = ToBool(#name, kv[#name])
#define GET_S(name) if (has(#name)) name = kv[#name]
  GET_S(objective); GET_S(boosting); GET_S(tree_learner);
  GET_I(num_iterations); GET_D(learning_rate); GET_I(num_leaves); GET_I(max_depth); GET_I(max_bin);
  GET_I(min_data_in_bin); GET_I(bin_construct_sample_cnt); GET_I(min_data_in_leaf); GET_D(min_sum_hessian_in_leaf);
  GET_D(lambda_l1); GET_D(lambda_l2); GET_D(min_gain_to_split); GET_D(max_delta_step); GET_B(boost_from_average);
  GET_D(sigmoid); GET_B(is_unbalance); GET_D(scale_pos_weight); GET_B(use_missing); GET_B(zero_as_missing);
  GET_B(feature_pre_filter); GET_B(use_quantized_grad); GET_I(num_grad_quant_bins); GET_B(stochastic_rounding);
  GET_B(quant_train_renew_leaf); GET_I(seed); GET_I(num_machines); GET_I(num_threads); GET_I(verbosity);
  GET_B(lambdarank_norm); GET_I(num_class);
  if (has("seed")) {      // upstream Config::Set: one LCG stream seeded with `seed` hands out the specific seeds, in this order
    HostRandom r(seed);
    r.NextShort(0, 32767);                          // data_random_seed (the host-side sample has its own)
    bagging_seed = r.NextShort(0, 32767);
    drop_seed = r.NextShort(0, 32767);
    feature_fraction_seed = r.NextShort(0, 32767);
    objective_seed = r.NextShort(0, 32767);
  }
  GET_D(alpha); GET_D(fair_c); GET_D(poisson_max_delta_step); GET_D(tweedie_variance_power);
  GET_D(feature_fraction); GET_D(feature_fraction_bynode); GET_I(feature_fraction_seed);
  GET_D(bagging_fraction); GET_D(pos_bagging_fraction); GET_D(neg_bagging_fraction); GET_I(bagging_freq); GET_I(bagging_seed);
  GET_D(drop_rate); GET_D(skip_drop); GET_D(top_rate); GET_D(other_rate); GET_I(max_drop); GET_I(drop_seed); GET_I(objective_seed); GET_B(uniform_drop); GET_B(xgboost_dart_mode);
  // SynapseML emits the pre-3.0 key `max_position` (RankerTrainParams.scala:28).  It is honoured as an alias of
  // lambdarank_truncation_level unless that key is given explicitly (documented choice).
  if (has("lambdarank_truncation_level")) lambdarank_truncation_level = static_cast<int>(ToDouble("lambdarank_truncation_level", kv["lambdarank_truncation_level"]));
  else if (has("max_position")) lambdarank_truncation_level = static_cast<int>(ToDouble("max_position", kv["max_position"]));
  if (has("label_gain")) label_gain = ToList<double>("label_gain", kv["label_gain"]);
  if (has("eval_at")) eval_at = ToList<int>("eval_at", kv["eval_at"]);
  if (has("categorical_feature")) categorical_feature = ToList<int>("categorical_feature", kv["categorical_feature"]);
  if (has("max_bin_by_feature")) max_bin_by_feature = ToList<int>("max_bin_by_feature", kv["max_bin_by_feature"]);
  if (has("monotone_constraints")) monotone_constraints = ToList<int>("monotone_constraints", kv["monotone_constraints"]);
  if (has("monotone_constraints_method")) monotone_constraints_method = kv["monotone_constraints_method"];
  GET_D(monotone_penalty);
  GET_I(max_cat_to_onehot); GET_I(max_cat_threshold); GET_I(min_data_per_group); GET_D(cat_smooth); GET_D(cat_l2);
#undef GET_I
#undef GET_D
#undef GET_B
#undef GET_S
  static const char* const kL2[] = {"regression_l2", "l2", "mean_squared_error", "mse", "l2_root", "root_mean_squared_error", "rmse"};
  for (auto a : kL2)
    if (objective == a) objective = "regression";
  if (objective == "softmax") objective = "multiclass";
  if (objective == "l1" || objective == "mean_absolute_error" || objective == "mae") objective = "regression_l1";
  if (objective == "mean_absolute_percentage_error") objective = "mape";
  if (objective == "xentropy") objective = "cross_entropy";
  if (objective == "xendcg" || objective == "xe_ndcg" || objective == "xe_ndcg_mart" || objective == "xendcg_mart") objective = "rank_xendcg";
  if (boosting == "random_forest") boosting = "rf";
  if (boosting == "gbrt") boosting = "gbdt";
}

}  // namespace lgb
