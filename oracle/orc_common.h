// ORACLE — TEST INFRASTRUCTURE ONLY.  Never linked, imported or executed by the product path
// (synapseml_b200/); only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
// legs may use it.
//
// PARITY UNPINNED: the arithmetic of the LightGBM training hot path lives in the third-party Maven artifact
// com.microsoft.ml.lightgbm:lightgbmlib:3.3.510 (/root/reference/build.sbt:297), whose source is not in
// /root/reference and which cannot be built or run in this container (no JVM, no lightgbm wheel).  This file set
// restates the published upstream microsoft/LightGBM algorithm (v3.3.x -> v4.0 line) from
// its specification, and anchors to the reference's own call sites (cited per function).
//
// Common constants, numeric helpers and the "k=v k=v" parameter-string parser
// (reference emits that string at lightgbm/.../params/BaseTrainParams.scala:25-40 and
//  lightgbm/.../LightGBMBase.scala:374-386; parsed natively by upstream Config::Str2Map).
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <map>
#include <set>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace orc {

typedef float score_t;    // upstream score_t
typedef float label_t;    // upstream label_t
typedef int32_t data_size_t;

static const double kZeroThreshold = 1e-35f;   // float literal widened, as upstream
static const double kEpsilon = 1e-15f;
static const double kSparseThreshold = 0.7;
static const double kMinScore = -std::numeric_limits<double>::infinity();

enum MissingType { MISSING_NONE = 0, MISSING_ZERO = 1, MISSING_NAN = 2 };

inline int RoundInt(double x) { return static_cast<int>(x + 0.5f); }
inline double UpperBoundOf(double a) { return std::nextafter(a, std::numeric_limits<double>::infinity()); }
inline bool EqualOrdered(double a, double b) { return b <= UpperBoundOf(a); }
inline int Sign(double x) { return (x > 0.0) - (x < 0.0); }

// Software exp shared (as an algorithm, not as source) with the CUDA gradient kernel so that float32 gradients are
// bit-identical on both sides: Cody-Waite reduction by ln2 (hi/lo), degree-13 Taylor evaluated by Horner with explicit
// fma, exact 2^k scaling.  Upstream calls std::exp here (binary_objective.hpp GetGradients); this differs from glibc exp
// by <= 1 ulp in double, which the float32 cast almost always absorbs (tests quantify the residue).
inline double SoftExp(double x) {
  if (x > 708.0) x = 708.0;
  if (x < -708.0) x = -708.0;
  const double kLog2e = 1.4426950408889634074;
  const double kLn2Hi = 6.93147180369123816490e-01;
  const double kLn2Lo = 1.90821492927058770002e-10;
  double kf = std::rint(x * kLog2e);
  double r = std::fma(-kf, kLn2Hi, x);
  r = std::fma(-kf, kLn2Lo, r);
  double p = 1.0 / 6227020800.0;
  p = std::fma(p, r, 1.0 / 479001600.0);
  p = std::fma(p, r, 1.0 / 39916800.0);
  p = std::fma(p, r, 1.0 / 3628800.0);
  p = std::fma(p, r, 1.0 / 362880.0);
  p = std::fma(p, r, 1.0 / 40320.0);
  p = std::fma(p, r, 1.0 / 5040.0);
  p = std::fma(p, r, 1.0 / 720.0);
  p = std::fma(p, r, 1.0 / 120.0);
  p = std::fma(p, r, 1.0 / 24.0);
  p = std::fma(p, r, 1.0 / 6.0);
  p = std::fma(p, r, 0.5);
  p = std::fma(p, r, 1.0);
  p = std::fma(p, r, 1.0);
  int k = static_cast<int>(kf);
  uint64_t bits = static_cast<uint64_t>(k + 1023) << 52;
  double s;
  std::memcpy(&s, &bits, sizeof(s));
  return p * s;
}

// Software log2 for x > 0 (normal range), same role as SoftExp: used by the lambdarank normalisation
// log2(1 + sum_lambdas) / sum_lambdas (rank_objective.hpp) so CPU and GPU agree bit for bit.
// x = m * 2^e with m in [sqrt(1/2), sqrt(2)); log(m) = 2 atanh(s), s = (m-1)/(m+1), odd series to s^23.
inline double SoftLog2(double x) {
  uint64_t bits;
  std::memcpy(&bits, &x, sizeof(bits));
  int e = static_cast<int>((bits >> 52) & 0x7ff) - 1023;
  bits = (bits & 0x000fffffffffffffULL) | 0x3ff0000000000000ULL;
  double m;
  std::memcpy(&m, &bits, sizeof(m));
  if (m > 1.4142135623730951) { m = m * 0.5; e += 1; }
  double s = (m - 1.0) / (m + 1.0);
  double z = s * s;
  double p = 1.0 / 23.0;
  p = std::fma(p, z, 1.0 / 21.0);
  p = std::fma(p, z, 1.0 / 19.0);
  p = std::fma(p, z, 1.0 / 17.0);
  p = std::fma(p, z, 1.0 / 15.0);
  p = std::fma(p, z, 1.0 / 13.0);
  p = std::fma(p, z, 1.0 / 11.0);
  p = std::fma(p, z, 1.0 / 9.0);
  p = std::fma(p, z, 1.0 / 7.0);
  p = std::fma(p, z, 1.0 / 5.0);
  p = std::fma(p, z, 1.0 / 3.0);
  p = std::fma(p, z, 1.0);
  double ln_m = 2.0 * s * p;
  const double kInvLn2 = 1.4426950408889634074;
  return std::fma(ln_m, kInvLn2, static_cast<double>(e));
}

// Counter-based uniform [0,1) used for stochastic rounding of quantized gradients.  Upstream draws from per-OpenMP-block
// std::mt19937 streams (gradient_discretizer.cpp), which depends on the thread count; this project defines its own
// stream keyed by (seed, iteration, global row, which) — a documented deviation, identical in oracle and kernel.
inline float HashUniform(uint32_t seed, uint32_t iter, uint32_t row, uint32_t which) {
  uint32_t h = seed * 0x9E3779B1u + iter * 0x85EBCA77u + row * 0xC2B2AE3Du + which * 0x27D4EB2Fu + 0x165667B1u;
  h ^= h >> 15; h *= 0x2C1B3C6Du; h ^= h >> 12; h *= 0x297A2D39u; h ^= h >> 15;
  return static_cast<float>(h >> 8) * (1.0f / 16777216.0f);
}

// include/LightGBM/utils/random.h (upstream): x = 214013 x + 2531011 (mod 2^32); 15-bit draws for NextShort / NextFloat,
// 31-bit draws for NextInt.
struct Random {
  unsigned int x = 123456789u;
  Random() {}
  explicit Random(int seed) : x(static_cast<unsigned int>(seed)) {}
  int RandInt16() { x = 214013u * x + 2531011u; return static_cast<int>((x >> 16) & 0x7FFF); }
  int RandInt32() { x = 214013u * x + 2531011u; return static_cast<int>(x & 0x7FFFFFFF); }
  int NextShort(int lo, int hi) { return RandInt16() % (hi - lo) + lo; }
  int NextInt(int lo, int hi) { return RandInt32() % (hi - lo) + lo; }
  float NextFloat() { return static_cast<float>(RandInt16()) / 32768.0f; }
  // Random::Sample(N, K): K distinct values of [0, N), ascending
  std::vector<int> Sample(int N, int K) {
    std::vector<int> ret;
    if (K > N || K <= 0) return ret;
    if (K == N) { for (int i = 0; i < N; ++i) ret.push_back(i); return ret; }
    if (K > 1 && K > (N / std::log2(static_cast<double>(K)))) {
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
      for (int v : chosen) ret.push_back(v);
    }
    return ret;
  }
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
  double lambda_l1 = 0.0;
  double lambda_l2 = 0.0;
  double min_gain_to_split = 0.0;
  double max_delta_step = 0.0;
  bool boost_from_average = true;
  double sigmoid = 1.0;
  bool is_unbalance = false;
  double scale_pos_weight = 1.0;
  bool use_missing = true;
  bool zero_as_missing = false;
  bool feature_pre_filter = true;
  bool use_quantized_grad = false;
  int num_grad_quant_bins = 4;
  int num_class = 1;
  bool stochastic_rounding = true;
  bool quant_train_renew_leaf = false;
  int seed = 0;
  // column / row sampling (upstream config.h defaults); `seed` derives the specific seeds (Config::Set, upstream src/io/config.cpp)
  // regression objectives beyond L2 (upstream config.h defaults)
  double alpha = 0.9, fair_c = 1.0, poisson_max_delta_step = 0.7, tweedie_variance_power = 1.5;
  double feature_fraction = 1.0, feature_fraction_bynode = 1.0;
  int feature_fraction_seed = 2;
  double bagging_fraction = 1.0, pos_bagging_fraction = 1.0, neg_bagging_fraction = 1.0;
  int bagging_freq = 0, bagging_seed = 3;
  // boosting = dart / goss (upstream config.h defaults)
  double drop_rate = 0.1, skip_drop = 0.5, top_rate = 0.2, other_rate = 0.1;
  int max_drop = 50, drop_seed = 4;
  int objective_seed = 5;
  bool uniform_drop = false, xgboost_dart_mode = false;
  int num_machines = 1;
  int num_threads = 0;
  int verbosity = 1;
  int lambdarank_truncation_level = 30;
  bool lambdarank_norm = true;
  std::vector<double> label_gain;
  std::vector<int> eval_at;
  // categorical features (upstream config.h defaults): column indexes, and the knobs of the categorical split search
  std::vector<int> categorical_feature;
  std::vector<int> max_bin_by_feature;
  std::vector<int> monotone_constraints;      // per column: +1 / -1 / 0
  std::string monotone_constraints_method = "basic";
  double monotone_penalty = 0.0;
  int max_cat_to_onehot = 4, max_cat_threshold = 32, min_data_per_group = 100;
  double cat_smooth = 10.0, cat_l2 = 10.0;
  std::map<std::string, std::string> raw;   // everything seen, canonical key -> value (first writer wins)

  static std::string Canon(const std::string& k) {
    static const std::map<std::string, std::string> alias = {
        {"objective_type", "objective"}, {"app", "objective"}, {"application", "objective"}, {"loss", "objective"},
        {"boosting_type", "boosting"}, {"boost", "boosting"},
        {"tree", "tree_learner"}, {"tree_type", "tree_learner"}, {"tree_learner_type", "tree_learner"},
        {"num_iteration", "num_iterations"}, {"n_iter", "num_iterations"}, {"num_tree", "num_iterations"},
        {"num_trees", "num_iterations"}, {"num_round", "num_iterations"}, {"num_rounds", "num_iterations"},
        {"nrounds", "num_iterations"}, {"num_boost_round", "num_iterations"}, {"n_estimators", "num_iterations"},
        {"max_iter", "num_iterations"},
        {"shrinkage_rate", "learning_rate"}, {"eta", "learning_rate"},
        {"num_leaf", "num_leaves"}, {"max_leaves", "num_leaves"}, {"max_leaf", "num_leaves"},
        {"max_leaf_nodes", "num_leaves"},
        {"min_data_per_leaf", "min_data_in_leaf"}, {"min_data", "min_data_in_leaf"},
        {"min_child_samples", "min_data_in_leaf"}, {"min_samples_leaf", "min_data_in_leaf"},
        {"min_sum_hessian_per_leaf", "min_sum_hessian_in_leaf"}, {"min_sum_hessian", "min_sum_hessian_in_leaf"},
        {"min_hessian", "min_sum_hessian_in_leaf"}, {"min_child_weight", "min_sum_hessian_in_leaf"},
        {"reg_alpha", "lambda_l1"}, {"l1_regularization", "lambda_l1"},
        {"reg_lambda", "lambda_l2"}, {"lambda", "lambda_l2"}, {"l2_regularization", "lambda_l2"},
        {"min_split_gain", "min_gain_to_split"},
        {"max_bins", "max_bin"},
        {"subsample_for_bin", "bin_construct_sample_cnt"},
        {"unbalance", "is_unbalance"}, {"unbalanced_sets", "is_unbalance"},
        {"num_thread", "num_threads"}, {"nthread", "num_threads"}, {"nthreads", "num_threads"}, {"n_jobs", "num_threads"},
        {"random_seed", "seed"}, {"random_state", "seed"},
        {"sub_feature_bynode", "feature_fraction_bynode"}, {"colsample_bynode", "feature_fraction_bynode"},
        {"sub_feature", "feature_fraction"}, {"colsample_bytree", "feature_fraction"},
        {"sub_row", "bagging_fraction"}, {"subsample", "bagging_fraction"}, {"bagging", "bagging_fraction"},
        {"subsample_freq", "bagging_freq"}, {"bagging_fraction_seed", "bagging_seed"},
        {"num_machine", "num_machines"},
        {"verbose", "verbosity"},
        {"ndcg_eval_at", "eval_at"}, {"ndcg_at", "eval_at"},
        {"mc", "monotone_constraints"}, {"monotone_constraint", "monotone_constraints"}, {"monotonic_cst", "monotone_constraints"},
        {"monotone_constraining_method", "monotone_constraints_method"}, {"mc_method", "monotone_constraints_method"},
        {"monotone_splits_penalty", "monotone_penalty"}, {"ms_penalty", "monotone_penalty"}, {"mc_penalty", "monotone_penalty"},
        {"cat_feature", "categorical_feature"}, {"categorical_column", "categorical_feature"}, {"cat_column", "categorical_feature"},
        {"categorical_features", "categorical_feature"},
    };
    auto it = alias.find(k);
    return it == alias.end() ? k : it->second;
  }

  static bool ParseBool(const std::string& v) {
    std::string s;
    for (char c : v) s.push_back(static_cast<char>(std::tolower(c)));
    if (s == "true" || s == "1" || s == "+") return true;
    if (s == "false" || s == "0" || s == "-") return false;
    throw std::runtime_error("Cannot parse bool parameter value: " + v);
  }

  template <typename T>
  static std::vector<T> ParseList(const std::string& v) {
    std::vector<T> out;
    std::stringstream ss(v);
    std::string tok;
    while (std::getline(ss, tok, ',')) {
      if (!tok.empty()) out.push_back(static_cast<T>(std::stod(tok)));
    }
    return out;
  }

  void Set(const std::string& params) {
    std::stringstream ss(params);
    std::string tok;
    std::map<std::string, std::string> kv;
    while (ss >> tok) {
      size_t eq = tok.find('=');
      if (eq == std::string::npos) continue;
      std::string k = Canon(tok.substr(0, eq));
      std::string v = tok.substr(eq + 1);
      if (kv.count(k) == 0) kv[k] = v;   // first writer wins (upstream KV2Map keeps the first, warns on the rest)
    }
    for (auto& p : kv) raw[p.first] = p.second;
    auto has = [&](const char* k) { return kv.count(k) > 0; };
    auto geti = [&](const char* k, int* dst) { if (has(k)) *dst = static_cast<int>(std::stod(kv[k])); };
    auto getd = [&](const char* k, double* dst) { if (has(k)) *dst = std::stod(kv[k]); };
    auto getb = [&](const char* k, bool* dst) { if (has(k)) *dst = ParseBool(kv[k]); };
    if (has("objective")) objective = kv["objective"];
    if (has("boosting")) boosting = kv["boosting"];
    if (has("tree_learner")) tree_learner = kv["tree_learner"];
    geti("num_iterations", &num_iterations);
    getd("learning_rate", &learning_rate);
    geti("num_leaves", &num_leaves);
    geti("max_depth", &max_depth);
    geti("max_bin", &max_bin);
    geti("min_data_in_bin", &min_data_in_bin);
    geti("bin_construct_sample_cnt", &bin_construct_sample_cnt);
    geti("min_data_in_leaf", &min_data_in_leaf);
    getd("min_sum_hessian_in_leaf", &min_sum_hessian_in_leaf);
    getd("lambda_l1", &lambda_l1);
    getd("lambda_l2", &lambda_l2);
    getd("min_gain_to_split", &min_gain_to_split);
    getd("max_delta_step", &max_delta_step);
    getb("boost_from_average", &boost_from_average);
    getd("sigmoid", &sigmoid);
    getb("is_unbalance", &is_unbalance);
    getd("scale_pos_weight", &scale_pos_weight);
    getb("use_missing", &use_missing);
    getb("zero_as_missing", &zero_as_missing);
    getb("feature_pre_filter", &feature_pre_filter);
    getb("use_quantized_grad", &use_quantized_grad);
    geti("num_grad_quant_bins", &num_grad_quant_bins);
    geti("num_class", &num_class);
    getb("stochastic_rounding", &stochastic_rounding);
    getb("quant_train_renew_leaf", &quant_train_renew_leaf);
    geti("seed", &seed);
    if (has("seed")) {     // upstream: Random rand(seed); the specific seeds are NextShort(0, int16 max) draws in declaration order
      Random r(seed);
      const int int_max = 32767;
      r.NextShort(0, int_max);                       // data_random_seed
      bagging_seed = r.NextShort(0, int_max);
      drop_seed = r.NextShort(0, int_max);
      feature_fraction_seed = r.NextShort(0, int_max);
      objective_seed = r.NextShort(0, int_max);
    }
    getd("alpha", &alpha);
    getd("fair_c", &fair_c);
    getd("poisson_max_delta_step", &poisson_max_delta_step);
    getd("tweedie_variance_power", &tweedie_variance_power);
    getd("feature_fraction", &feature_fraction);
    getd("feature_fraction_bynode", &feature_fraction_bynode);
    geti("feature_fraction_seed", &feature_fraction_seed);
    getd("bagging_fraction", &bagging_fraction);
    getd("pos_bagging_fraction", &pos_bagging_fraction); getd("neg_bagging_fraction", &neg_bagging_fraction);
    geti("bagging_freq", &bagging_freq);
    geti("bagging_seed", &bagging_seed);
    getd("drop_rate", &drop_rate); getd("skip_drop", &skip_drop); getd("top_rate", &top_rate); getd("other_rate", &other_rate);
    geti("objective_seed", &objective_seed);
    geti("max_drop", &max_drop); geti("drop_seed", &drop_seed); getb("uniform_drop", &uniform_drop); getb("xgboost_dart_mode", &xgboost_dart_mode);
    if (boosting == "random_forest") boosting = "rf";
    if (boosting == "gbrt") boosting = "gbdt";
    geti("num_machines", &num_machines);
    geti("num_threads", &num_threads);
    geti("verbosity", &verbosity);
    // SynapseML still sends the pre-3.0 key max_position (RankerTrainParams.scala:28); honoured as an alias of
    // lambdarank_truncation_level unless the latter is given explicitly (SURVEY.md F10 — a documented choice).
    if (has("lambdarank_truncation_level")) geti("lambdarank_truncation_level", &lambdarank_truncation_level);
    else if (has("max_position")) geti("max_position", &lambdarank_truncation_level);
    getb("lambdarank_norm", &lambdarank_norm);
    if (has("label_gain")) label_gain = ParseList<double>(kv["label_gain"]);
    if (has("eval_at")) eval_at = ParseList<int>(kv["eval_at"]);
    if (has("categorical_feature")) categorical_feature = ParseList<int>(kv["categorical_feature"]);
    if (has("max_bin_by_feature")) max_bin_by_feature = ParseList<int>(kv["max_bin_by_feature"]);
    if (has("monotone_constraints")) monotone_constraints = ParseList<int>(kv["monotone_constraints"]);
    if (has("monotone_constraints_method")) monotone_constraints_method = kv["monotone_constraints_method"];
    getd("monotone_penalty", &monotone_penalty);
    geti("max_cat_to_onehot", &max_cat_to_onehot); geti("max_cat_threshold", &max_cat_threshold); geti("min_data_per_group", &min_data_per_group);
    getd("cat_smooth", &cat_smooth); getd("cat_l2", &cat_l2);
    // objective aliases
    if (objective == "regression_l2" || objective == "l2" || objective == "mean_squared_error" || objective == "mse" ||
        objective == "l2_root" || objective == "root_mean_squared_error" || objective == "rmse")
      objective = "regression";
    if (objective == "softmax") objective = "multiclass";
    if (objective == "l1" || objective == "mean_absolute_error" || objective == "mae") objective = "regression_l1";
    if (objective == "mean_absolute_percentage_error") objective = "mape";
    if (objective == "xentropy") objective = "cross_entropy";
    if (objective == "xendcg" || objective == "xe_ndcg" || objective == "xe_ndcg_mart" || objective == "xendcg_mart") objective = "rank_xendcg";
  }
};

}  // namespace orc
