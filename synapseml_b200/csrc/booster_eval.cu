// Validation data and evaluation metrics of the booster: LGBM_BoosterAddValidData /
// GetEvalCounts / GetEvalNames / GetEval as used by
// lightgbm/src/main/scala/com/microsoft/azure/synapse/ml/lightgbm/TrainUtils.scala:137-169 and booster/LightGBMBooster.scala:283-314.
// Scores of a validation set live on the device and are advanced by the same apply-tree kernel as the training scores
// (booster.cu); the metrics are device reductions (metric_kernels.cuh).  Metrics are local to the process.
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include <algorithm>
#include <cmath>
#include <sstream>

#include "booster.h"
#include "metric_kernels.cuh"

namespace lgb {

namespace {
constexpr int kNumSms = 148;
enum { kEvLogloss = 0, kEvError = 1, kEvL2 = 2, kEvRmse = 3, kEvAuc = 4, kEvNdcg = 5, kEvHuber = 6, kEvFair = 7, kEvPoisson = 8, kEvGamma = 9,
       kEvTweedie = 10, kEvMultiLogloss = 11, kEvMultiError = 12, kEvL1 = 13, kEvQuantile = 14, kEvMape = 15, kEvGammaDeviance = 16, kEvMap = 17, kEvXentropy = 18 };   // (Booster has its own kL2: distinct names)
}

void Booster::InitMetrics() {
  if (metrics_ready_) return;
  metrics_.clear();
  std::string spec;
  for (const char* key : {"metric", "metrics", "metric_types"}) {
    auto it = cfg_.raw.find(key);
    if (it != cfg_.raw.end()) { spec = it->second; break; }
  }
  std::vector<std::string> names;
  {
    std::string t;
    for (char ch : spec + ",") {
      if (ch == ',') { if (!t.empty()) names.push_back(t); t.clear(); }
      else if (ch != ' ' && ch != '"') t.push_back(ch);
    }
  }
  if (names.empty() && obj_ != kCustom)      // the objective's default; a custom objective has none
    names.push_back(obj_ == kBinary ? "binary_logloss" : obj_ == kL2 ? "l2" : (obj_ == kLambdarank || obj_ == kXendcg) ? "ndcg" : (obj_ == kMulticlass || obj_ == kMulticlassOva) ? "multi_logloss"
                    : cfg_.objective);
  for (const std::string& m : names) {
    if (m == "None" || m == "none" || m == "null" || m == "na" || m == "custom") continue;
    if (m == "binary_logloss" || m == "binary") metrics_.push_back({kEvLogloss, "binary_logloss"});
    else if (m == "binary_error") metrics_.push_back({kEvError, "binary_error"});
    else if (m == "auc") metrics_.push_back({kEvAuc, "auc"});
    else if (m == "l2" || m == "mse" || m == "mean_squared_error" || m == "regression" || m == "regression_l2") metrics_.push_back({kEvL2, "l2"});
    else if (m == "rmse" || m == "root_mean_squared_error" || m == "l2_root") metrics_.push_back({kEvRmse, "rmse"});
    else if (m == "huber") metrics_.push_back({kEvHuber, "huber"});
    else if (m == "fair") metrics_.push_back({kEvFair, "fair"});
    else if (m == "poisson") metrics_.push_back({kEvPoisson, "poisson"});
    else if (m == "gamma") metrics_.push_back({kEvGamma, "gamma"});
    else if (m == "tweedie") metrics_.push_back({kEvTweedie, "tweedie"});
    else if (m == "l1" || m == "mae" || m == "mean_absolute_error" || m == "regression_l1") metrics_.push_back({kEvL1, "l1"});
    else if (m == "quantile") metrics_.push_back({kEvQuantile, "quantile"});
    else if (m == "mape" || m == "mean_absolute_percentage_error") metrics_.push_back({kEvMape, "mape"});
    else if (m == "gamma_deviance" || m == "gamma-deviance") metrics_.push_back({kEvGammaDeviance, "gamma_deviance"});
    else if (m == "cross_entropy" || m == "xentropy") metrics_.push_back({kEvXentropy, "cross_entropy"});
    else if (m == "multi_logloss" || m == "multiclass" || m == "softmax" || m == "multiclassova" || m == "multiclass_ova" || m == "ova" || m == "ovr") metrics_.push_back({kEvMultiLogloss, "multi_logloss"});
    else if (m == "multi_error") metrics_.push_back({kEvMultiError, "multi_error"});
    else if (m == "ndcg" || m == "lambdarank" || m == "rank_xendcg" || m == "xendcg") {
      std::vector<int> at = cfg_.eval_at;
      if (at.empty()) at = {1, 2, 3, 4, 5};
      std::sort(at.begin(), at.end());
      LGB_CHECK(static_cast<int>(at.size()) <= kMaxEvalAt && at.front() > 0, "eval_at: between 1 and %d positive positions are supported", kMaxEvalAt);
      for (int k : at) metrics_.push_back({kEvNdcg, "ndcg@" + std::to_string(k)});
    } else if (m == "map" || m == "mean_average_precision") {
      std::vector<int> at = cfg_.eval_at;
      if (at.empty()) at = {1, 2, 3, 4, 5};
      std::sort(at.begin(), at.end());
      LGB_CHECK(static_cast<int>(at.size()) <= kMaxEvalAt && at.front() > 0, "eval_at: between 1 and %d positive positions are supported", kMaxEvalAt);
      for (int k : at) metrics_.push_back({kEvMap, "map@" + std::to_string(k)});
    } else {
      LGB_CHECK(false, "metric=%s is not supported by this library (binary_logloss, binary_error, auc, l2, rmse, ndcg, map, ...)", m.c_str());
    }
  }
  metrics_ready_ = true;
}

std::vector<std::string> Booster::EvalNames() {
  LGB_CHECK(!predict_only_, "a booster loaded from a model string has no metrics");
  InitMetrics();
  std::vector<std::string> out;
  for (auto& m : metrics_) out.push_back(m.name);
  return out;
}

void Booster::AddValidData(Dataset* valid) {
  LGB_CHECK(!predict_only_, "a booster loaded from a model string cannot take validation data");
  LGB_CHECK(valid != nullptr && valid->finished, "AddValidData: the validation dataset is missing or not finished");
  LGB_CHECK(valid->num_total_features == ds_->num_total_features && valid->num_features() == ds_->num_features(),
            "AddValidData: the validation dataset was not created from the training dataset's bin mappers");
  LGB_CHECK(valid->device == device_, "AddValidData: the validation dataset lives on another GPU");
  LGB_CHECK(iter_ == 0 && static_cast<int>(models_.size()) == num_init_iteration_, "AddValidData must be called before the first iteration");
  LGB_CUDA(cudaSetDevice(device_));
  valids_.emplace_back();
  ValidSet& v = valids_.back();
  v.ds = valid;
  v.score.AllocZero(std::max<int64_t>(valid->num_data, 1) * K_);      // class-major [K][n]
  if (valid->has_init_score)
    LGB_CUDA(cudaMemcpy(v.score.p, valid->d_init_score.p, sizeof(double) * valid->num_data * K_, cudaMemcpyDeviceToDevice));
}

// max DCG per query and position, and query weights (mean row weight), from the host copies of labels / weights
void Booster::PrepareNdcg(ValidSet* v, const Dataset* ds) {
  if (v->ndcg_ready) return;
  LGB_CHECK(ds->query_boundaries.size() >= 2, "ndcg needs query information on the evaluated dataset");
  std::vector<int> at;
  for (auto& m : metrics_) if (m.kind == kEvNdcg) at.push_back(atoi(m.name.c_str() + 5));
  const int K = static_cast<int>(at.size());
  std::vector<double> label_gain = cfg_.label_gain;
  if (label_gain.empty()) for (int i = 0; i < 31; ++i) label_gain.push_back(static_cast<double>((1LL << i) - 1));
  const int nq = static_cast<int>(ds->query_boundaries.size()) - 1;
  std::vector<double> inv(static_cast<size_t>(nq) * K), qw;
  std::vector<int> cnt(label_gain.size());
  for (int q = 0; q < nq; ++q) {
    const int b = ds->query_boundaries[q], e = ds->query_boundaries[q + 1];
    std::fill(cnt.begin(), cnt.end(), 0);
    for (int i = b; i < e; ++i) {
      const int l = static_cast<int>(ds->label[i]);
      LGB_CHECK(l >= 0 && l < static_cast<int>(label_gain.size()), "ndcg: label %d is outside the label_gain table", l);
      ++cnt[l];
    }
    // CalMaxDCG: the best ordering puts the highest labels first
    double dcg = 0.0;
    int top = static_cast<int>(label_gain.size()) - 1, pos = 0;
    for (int j = 0; j < K; ++j) {
      const int upto = std::min(at[j], e - b);
      for (; pos < upto; ++pos) {
        while (top > 0 && cnt[top] <= 0) --top;
        if (top < 0 || cnt[top] <= 0) break;
        dcg += label_gain[top] / std::log2(2.0 + pos);
        --cnt[top];
      }
      inv[static_cast<size_t>(q) * K + j] = dcg > 0.0 ? 1.0 / dcg : -1.0;
    }
  }
  v->inv_max_dcg.Alloc(inv.size());
  LGB_CUDA(cudaMemcpy(v->inv_max_dcg.p, inv.data(), sizeof(double) * inv.size(), cudaMemcpyHostToDevice));
  if (ds->has_weight) {
    qw.resize(nq);
    for (int q = 0; q < nq; ++q) {
      const int b = ds->query_boundaries[q], e = ds->query_boundaries[q + 1];
      double s = 0.0;
      for (int i = b; i < e; ++i) s += ds->weight[i];
      qw[q] = e > b ? s / (e - b) : 0.0;
    }
    v->qweight.Alloc(nq);
    LGB_CUDA(cudaMemcpy(v->qweight.p, qw.data(), sizeof(double) * nq, cudaMemcpyHostToDevice));
  }
  if (d_label_gain_.p == nullptr) {
    d_label_gain_.Alloc(label_gain.size());
    LGB_CUDA(cudaMemcpy(d_label_gain_.p, label_gain.data(), sizeof(double) * label_gain.size(), cudaMemcpyHostToDevice));
  }
  v->ndcg_ready = true;
}

void Booster::Eval(int data_idx, std::vector<double>* out) {
  LGB_CHECK(!predict_only_, "a booster loaded from a model string has no metrics");
  InitMetrics();
  LGB_CHECK(data_idx >= 0 && data_idx <= static_cast<int>(valids_.size()), "GetEval: data_idx %d is out of range (%d validation sets)",
            data_idx, static_cast<int>(valids_.size()));
  LGB_CUDA(cudaSetDevice(device_));
  ValidSet* vs = data_idx == 0 ? &train_eval_ : &valids_[data_idx - 1];
  const Dataset* ds = data_idx == 0 ? ds_ : vs->ds;
  const double* score = data_idx == 0 ? d_score_.p : vs->score.p;
  const long long n = ds->num_data;
  const float* w = ds->has_weight ? ds->d_weight.p : nullptr;
  out->clear();
  d_mpart_.Reserve(kMetricBlocks);
  d_mout_.Reserve(kMaxEvalAt + 2);
  double h[kMaxEvalAt + 2];
  bool ndcg_done = false, map_done = false;
  for (const MetricSpec& m : metrics_) {
    if (m.kind == kEvMultiLogloss || m.kind == kEvMultiError) {
      LGB_CHECK(obj_ == kMulticlass || obj_ == kMulticlassOva, "metric=%s needs objective=multiclass / multiclassova", m.name.c_str());
      k_metric_multi<<<kMetricBlocks, kMetricThreads, 0, stream_>>>(score, ds->d_label.p, w, n, K_, m.kind == kEvMultiLogloss ? 0 : 1, d_mpart_.p,
                                                                    obj_ == kMulticlassOva ? cfg_.sigmoid : 0.0);
      k_metric_final<<<1, kMetricThreads, 0, stream_>>>(d_mpart_.p, kMetricBlocks, d_mout_.p);
      LGB_CUDA(cudaMemcpyAsync(h, d_mout_.p, sizeof(double) * 2, cudaMemcpyDeviceToHost, stream_));
      LGB_CUDA(cudaStreamSynchronize(stream_));
      out->push_back(h[0] / h[1]);
      continue;
    }
    LGB_CHECK(obj_ != kMulticlass && obj_ != kMulticlassOva, "metric=%s is not defined for multiclass objectives (multi_logloss, multi_error)", m.name.c_str());
    if (m.kind == kEvLogloss || m.kind == kEvError || m.kind == kEvL2 || m.kind == kEvRmse || (m.kind >= kEvHuber && m.kind != kEvMap)) {
      const int kind = m.kind == kEvLogloss ? kMetricBinaryLogloss : m.kind == kEvError ? kMetricBinaryError
                       : m.kind >= kEvL1 ? kMetricL1 + (m.kind - kEvL1) : m.kind >= kEvHuber ? kMetricHuber + (m.kind - kEvHuber) : kMetricL2;
      const double p0 = (m.kind == kEvHuber || m.kind == kEvQuantile) ? cfg_.alpha : m.kind == kEvFair ? cfg_.fair_c : cfg_.tweedie_variance_power;
      k_metric_point<<<kMetricBlocks, kMetricThreads, 0, stream_>>>(score, ds->d_label.p, w, n, kind, cfg_.sigmoid, d_mpart_.p,
                                                                    LogLink() ? 1 : 0, p0);
      k_metric_final<<<1, kMetricThreads, 0, stream_>>>(d_mpart_.p, kMetricBlocks, d_mout_.p);
      LGB_CUDA(cudaMemcpyAsync(h, d_mout_.p, sizeof(double) * 2, cudaMemcpyDeviceToHost, stream_));
      LGB_CUDA(cudaStreamSynchronize(stream_));
      const double v = h[0] / h[1];
      out->push_back(m.kind == kEvRmse ? std::sqrt(v) : m.kind == kEvGammaDeviance ? h[0] * 2 : v);      // GammaDevianceMetric::AverageLoss: sum_loss * 2
    } else if (m.kind == kEvAuc) {
      const int ni = static_cast<int>(n);
      d_auc_keys_[0].Reserve(n); d_auc_keys_[1].Reserve(n); d_auc_idx_[0].Reserve(n); d_auc_idx_[1].Reserve(n);
      d_auc_pos_.Reserve(n); d_auc_neg_.Reserve(n); d_auc_prefix_.Reserve(n);
      k_iota<<<std::min(CeilDiv(ni, 256), kNumSms * 8), 256, 0, stream_>>>(d_auc_idx_[0].p, ni);
      LGB_CUDA(cudaMemcpyAsync(d_auc_keys_[0].p, score, sizeof(double) * n, cudaMemcpyDeviceToDevice, stream_));
      size_t tmp = 0, tmp2 = 0;
      LGB_CUDA(cub::DeviceRadixSort::SortPairsDescending(nullptr, tmp, d_auc_keys_[0].p, d_auc_keys_[1].p, d_auc_idx_[0].p, d_auc_idx_[1].p, ni,
                                                         0, 64, stream_));
      LGB_CUDA(cub::DeviceScan::InclusiveSum(nullptr, tmp2, d_auc_pos_.p, d_auc_prefix_.p, ni, stream_));
      d_cub_tmp_.Reserve(std::max(tmp, tmp2));
      tmp = tmp2 = d_cub_tmp_.n;
      LGB_CUDA(cub::DeviceRadixSort::SortPairsDescending(d_cub_tmp_.p, tmp, d_auc_keys_[0].p, d_auc_keys_[1].p, d_auc_idx_[0].p, d_auc_idx_[1].p,
                                                         ni, 0, 64, stream_));
      k_auc_gather<<<kMetricBlocks, kMetricThreads, 0, stream_>>>(d_auc_idx_[1].p, ds->d_label.p, w, n, d_auc_pos_.p, d_auc_neg_.p);
      LGB_CUDA(cub::DeviceScan::InclusiveSum(d_cub_tmp_.p, tmp2, d_auc_pos_.p, d_auc_prefix_.p, ni, stream_));
      k_auc_accumulate<<<kMetricBlocks, kMetricThreads, 0, stream_>>>(d_auc_keys_[1].p, d_auc_prefix_.p, d_auc_neg_.p, n, d_mpart_.p);
      k_metric_final<<<1, kMetricThreads, 0, stream_>>>(d_mpart_.p, kMetricBlocks, d_mout_.p);
      double sum_pos = 0.0;
      LGB_CUDA(cudaMemcpyAsync(h, d_mout_.p, sizeof(double) * 2, cudaMemcpyDeviceToHost, stream_));
      LGB_CUDA(cudaMemcpyAsync(&sum_pos, d_auc_prefix_.p + (n - 1), sizeof(double), cudaMemcpyDeviceToHost, stream_));
      LGB_CUDA(cudaStreamSynchronize(stream_));
      const double accum = h[0], sum_neg = h[1];
      out->push_back((sum_pos > 0.0 && sum_neg > 0.0) ? accum / (sum_pos * sum_neg) : 1.0);
    } else if (m.kind == kEvNdcg) {
      if (ndcg_done) continue;      // all positions come out of one pass
      PrepareNdcg(vs, ds);
      NdcgArgs a;
      a.K = 0;
      for (auto& mm : metrics_) if (mm.kind == kEvNdcg) a.eval_at[a.K++] = atoi(mm.name.c_str() + 5);
      const int nq = static_cast<int>(ds->query_boundaries.size()) - 1;
      int max_cnt = 1;
      for (int q = 0; q < nq; ++q) max_cnt = std::max(max_cnt, ds->query_boundaries[q + 1] - ds->query_boundaries[q]);
      a.score = score; a.label = ds->d_label.p; a.qb = ds->d_query_boundaries.p; a.label_gain = d_label_gain_.p;
      a.num_gain = static_cast<int>(d_label_gain_.n); a.max_cnt = max_cnt;
      d_ndcg_.Reserve(static_cast<size_t>(nq) * a.K);
      a.dcg = d_ndcg_.p;
      const size_t smem = sizeof(double) * max_cnt;
      LGB_CHECK(smem <= 200 * 1024, "ndcg: a query with %d documents does not fit the kernel's shared memory", max_cnt);
      static size_t smem_set = 0;
      if (smem > smem_set) { LGB_CUDA(cudaFuncSetAttribute(k_ndcg, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem))); smem_set = smem; }
      k_ndcg<<<nq, kMetricThreads, smem, stream_>>>(a);
      k_ndcg_final<<<1, kMetricThreads, 0, stream_>>>(d_ndcg_.p, vs->inv_max_dcg.p, vs->qweight.p, nq, a.K, d_mout_.p);
      LGB_CUDA(cudaMemcpyAsync(h, d_mout_.p, sizeof(double) * (a.K + 1), cudaMemcpyDeviceToHost, stream_));
      LGB_CUDA(cudaStreamSynchronize(stream_));
      for (int j = 0; j < a.K; ++j) out->push_back(h[j] / h[a.K]);
      ndcg_done = true;
    } else if (m.kind == kEvMap) {
      if (map_done) continue;      // all positions come out of one pass
      PrepareNdcg(vs, ds);         // (query boundaries checked, query weights)
      NdcgArgs a;
      a.K = 0;
      for (auto& mm : metrics_) if (mm.kind == kEvMap) a.eval_at[a.K++] = atoi(mm.name.c_str() + 4);
      const int nq = static_cast<int>(ds->query_boundaries.size()) - 1;
      int max_cnt = 1;
      for (int q = 0; q < nq; ++q) max_cnt = std::max(max_cnt, ds->query_boundaries[q + 1] - ds->query_boundaries[q]);
      a.score = score; a.label = ds->d_label.p; a.qb = ds->d_query_boundaries.p; a.label_gain = nullptr; a.num_gain = 0; a.max_cnt = max_cnt;
      d_ndcg_.Reserve(static_cast<size_t>(nq) * a.K);
      a.dcg = d_ndcg_.p;
      const size_t smem = (sizeof(double) + sizeof(int)) * max_cnt;
      LGB_CHECK(smem <= 200 * 1024, "map: a query with %d documents does not fit the kernel's shared memory", max_cnt);
      static size_t smem_set = 0;
      if (smem > smem_set) { LGB_CUDA(cudaFuncSetAttribute(k_map, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem))); smem_set = smem; }
      k_map<<<nq, kMetricThreads, smem, stream_>>>(a);
      k_ndcg_final<<<1, kMetricThreads, 0, stream_>>>(d_ndcg_.p, nullptr, vs->qweight.p, nq, a.K, d_mout_.p);
      LGB_CUDA(cudaMemcpyAsync(h, d_mout_.p, sizeof(double) * (a.K + 1), cudaMemcpyDeviceToHost, stream_));
      LGB_CUDA(cudaStreamSynchronize(stream_));
      for (int j = 0; j < a.K; ++j) out->push_back(h[j] / h[a.K]);
      map_done = true;
    }
  }
  LGB_CUDA(cudaGetLastError());
}

}  // namespace lgb
