// Evaluation metrics on device-resident scores: what LGBM_BoosterGetEval returns to
// lightgbm/src/main/scala/com/microsoft/azure/synapse/ml/lightgbm/TrainUtils.scala:137-169 (early stopping) via
// booster/LightGBMBooster.scala:283-314.  Definitions follow upstream's metric classes (binary_metric.hpp, regression_metric.hpp,
// rank_metric.hpp / dcg_calculator.cpp): weighted means, kEpsilon = 1e-15f clamps, AUC with tied scores grouped, NDCG with
// gains 2^label - 1 and discounts 1 / log2(2 + rank), a query without relevant documents counting as 1.
// All reductions are two-level with a fixed order (per-CTA partials, then one CTA), so a metric is reproducible run to run.
#pragma once
#include <cstdint>

namespace lgb {

__global__ void k_iota(int* p, int n) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) p[i] = i;
}

constexpr int kMetricBlocks = 592;
constexpr int kMetricThreads = 256;

enum PointMetric { kMetricBinaryLogloss = 0, kMetricBinaryError = 1, kMetricL2 = 2, kMetricHuber = 3, kMetricFair = 4, kMetricPoisson = 5,
                   kMetricGamma = 6, kMetricTweedie = 7, kMetricL1 = 8, kMetricQuantile = 9, kMetricMape = 10, kMetricGammaDeviance = 11, kMetricXentropy = 13 };

__device__ __forceinline__ double BlockSumD(double v, double* s_w /*[8]*/) {
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = v;
  __syncthreads();
  double t = 0.0;
  for (int w = 0; w < kMetricThreads / 32; ++w) t += s_w[w];
  return t;
}

// partial[b] = (sum loss * w, sum w) over the rows of CTA b
__global__ void __launch_bounds__(kMetricThreads) k_metric_point(const double* __restrict__ score, const float* __restrict__ label,
                                                                  const float* __restrict__ weight, long long n, int kind, double sigmoid,
                                                                  double2* __restrict__ partial, int exp_link = 0, double p0 = 0.0) {
  __shared__ double s_w[8];
  const double kEps = (double)1e-15f;
  double sl = 0.0, sw = 0.0;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const double s = score[i];
    const float y = label[i];
    const double w = weight ? static_cast<double>(weight[i]) : 1.0;
    double loss;
    if (kind >= kMetricL2) {
      // regression_metric.hpp: the loss sees the objective's converted output (exp for the log-link objectives)
      const double t = exp_link ? exp(s) : s;
      const double lab = static_cast<double>(y);
      if (kind == kMetricL2) {
        const double d = t - lab;
        loss = d * d;
      } else if (kind == kMetricHuber) {          // p0 = alpha
        const double d = t - lab;
        loss = fabs(d) <= p0 ? 0.5 * d * d : p0 * (fabs(d) - 0.5 * p0);
      } else if (kind == kMetricFair) {           // p0 = fair_c
        const double x = fabs(t - lab);
        loss = p0 * x - p0 * p0 * log1p(x / p0);
      } else if (kind == kMetricPoisson) {
        const double sc = t < 1e-10 ? 1e-10 : t;
        loss = sc - lab * log(sc);
      } else if (kind == kMetricGamma) {          // psi = 1: -((label * theta - b) / a + c), theta = -1 / score
        const double theta = -1.0 / t;
        const double b = -((-theta) > 0.0 ? log(-theta) : -INFINITY);
        const double c = (lab > 0.0 ? log(lab) : -INFINITY) - (lab > 0.0 ? log(lab) : -INFINITY);
        loss = -((lab * theta - b) + c);
      } else if (kind == kMetricL1) {
        loss = fabs(t - lab);
      } else if (kind == kMetricQuantile) {       // p0 = alpha
        const double delta = lab - t;
        loss = delta < 0 ? (p0 - 1.0f) * delta : p0 * delta;
      } else if (kind == kMetricMape) {
        loss = fabs(lab - t) / fmax(1.0, fabs(lab));
      } else if (kind == kMetricXentropy) {       // CrossEntropyMetric: XentLoss(label, sigmoid(score)), logs clipped at 1e-12
        const double p = 1.0 / (1.0 + exp(-s));
        const double a = lab * (p > 1.0e-12 ? log(p) : log(1.0e-12));
        const double b = (1.0f - lab) * (1.0f - p > 1.0e-12 ? log(1.0f - p) : log(1.0e-12));
        loss = -(a + b);
      } else if (kind == kMetricGammaDeviance) {  // GammaDevianceMetric::LossOnPoint
        const double tmp = lab / (t + 1.0e-9);
        loss = tmp - (tmp > 0.0 ? log(tmp) : -INFINITY) - 1;
      } else {                                    // tweedie, p0 = rho
        const double sc = t < 1e-10 ? 1e-10 : t;
        const double a = lab * exp((1 - p0) * log(sc)) / (1 - p0);
        const double b = exp((2 - p0) * log(sc)) / (2 - p0);
        loss = -a + b;
      }
    } else {
      const double prob = 1.0 / (1.0 + exp(-sigmoid * s));
      if (kind == kMetricBinaryLogloss) {
        if (y <= 0.0f) loss = (1.0 - prob > kEps) ? -log(1.0 - prob) : -log(kEps);
        else loss = (prob > kEps) ? -log(prob) : -log(kEps);
      } else {
        loss = (prob <= 0.5) ? (y > 0.0f ? 1.0 : 0.0) : (y <= 0.0f ? 1.0 : 0.0);
      }
    }
    sl += loss * w; sw += w;
  }
  const double tl = BlockSumD(sl, s_w);
  const double tw = BlockSumD(sw, s_w);
  if (threadIdx.x == 0) partial[blockIdx.x] = make_double2(tl, tw);
}

// multi_logloss (kind 0) / multi_error (kind 1) on class-major scores [K][n] (upstream multiclass_metric.hpp): the loss sees the
// softmax probabilities; multi_error counts a row as wrong when another class scores at least as high as its label's
// (multi_error_top_k = 1).
__global__ void __launch_bounds__(kMetricThreads) k_metric_multi(const double* __restrict__ score, const float* __restrict__ label,
                                                                  const float* __restrict__ weight, long long n, int K, int kind,
                                                                  double2* __restrict__ partial, double ova_sigmoid = 0.0) {
  __shared__ double s_w[8];
  const double kEps = (double)1e-15f;
  double sl = 0.0, sw = 0.0;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int lab = static_cast<int>(label[i]);
    const double w = weight ? static_cast<double>(weight[i]) : 1.0;
    double wmax = score[i];
    for (int k = 1; k < K; ++k) wmax = fmax(wmax, score[static_cast<size_t>(k) * n + i]);
    double wsum = 0.0, pl = 0.0;
    int num_larger = 0;
    const double ref = score[static_cast<size_t>(lab) * n + i];
    for (int k = 0; k < K; ++k) {
      const double sc = score[static_cast<size_t>(k) * n + i];
      const double e = exp(sc - wmax);
      wsum += e;
      if (k == lab) pl = e;
      num_larger += sc >= ref;
    }
    double loss;
    // the loss sees the objective's ConvertOutput: softmax, or (multiclassova, ova_sigmoid > 0) one sigmoid per class
    if (kind == 0) { const double p = ova_sigmoid > 0.0 ? 1.0 / (1.0 + exp(-ova_sigmoid * ref)) : pl / wsum; loss = p > kEps ? -log(p) : -log(kEps); }
    else loss = num_larger > 1 ? 1.0 : 0.0;
    sl += loss * w; sw += w;
  }
  const double tl = BlockSumD(sl, s_w);
  const double tw = BlockSumD(sw, s_w);
  if (threadIdx.x == 0) partial[blockIdx.x] = make_double2(tl, tw);
}

// one CTA: out[0..1] = sums of the partials in index order
__global__ void __launch_bounds__(kMetricThreads) k_metric_final(const double2* __restrict__ partial, int n, double* __restrict__ out) {
  __shared__ double s_w[8];
  double a = 0.0, b = 0.0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) { a += partial[i].x; b += partial[i].y; }
  const double ta = BlockSumD(a, s_w);
  const double tb = BlockSumD(b, s_w);
  if (threadIdx.x == 0) { out[0] = ta; out[1] = tb; }
}

// ---- AUC.  Rows sorted by score, descending (keys) with their original index (order).
__global__ void k_auc_gather(const int* __restrict__ order, const float* __restrict__ label, const float* __restrict__ weight, long long n,
                             double* __restrict__ pos_w, double* __restrict__ neg_w) {
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int r = order[i];
    const double w = weight ? static_cast<double>(weight[r]) : 1.0;
    const bool pos = label[r] > 0.0f;
    pos_w[i] = pos ? w : 0.0;
    neg_w[i] = pos ? 0.0 : w;
  }
}
// pos_prefix: inclusive prefix sum of pos_w.  Every negative row contributes its weight times (positives ranked strictly
// above its score group + half of the positives inside the group) — upstream's accumulation, written per row.
__global__ void __launch_bounds__(kMetricThreads) k_auc_accumulate(const double* __restrict__ keys, const double* __restrict__ pos_prefix,
                                                                    const double* __restrict__ neg_w, long long n,
                                                                    double2* __restrict__ partial) {
  __shared__ double s_w[8];
  double acc = 0.0, negs = 0.0;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const double nw = neg_w[i];
    negs += nw;
    if (nw == 0.0) continue;
    const double k = keys[i];
    // keys are descending: first index with key == k, last index with key == k
    long long lo = 0, hi = i;
    while (lo < hi) { const long long m = (lo + hi) >> 1; if (keys[m] > k) lo = m + 1; else hi = m; }
    const long long first = lo;
    lo = i; hi = n - 1;
    while (lo < hi) { const long long m = (lo + hi + 1) >> 1; if (keys[m] < k) hi = m - 1; else lo = m; }
    const long long last = lo;
    const double before = first > 0 ? pos_prefix[first - 1] : 0.0;
    const double in_group = pos_prefix[last] - before;
    acc += nw * (before + 0.5 * in_group);
  }
  const double ta = BlockSumD(acc, s_w);
  const double tn = BlockSumD(negs, s_w);
  if (threadIdx.x == 0) partial[blockIdx.x] = make_double2(ta, tn);
}

// ---- NDCG@k.  One CTA per query; rank by counting (= std::stable_sort by score, descending).  out[q * K + j] = DCG@eval_at[j].
constexpr int kMaxEvalAt = 8;
struct NdcgArgs {
  const double* score; const float* label; const int* qb; const double* label_gain; int num_gain;
  int eval_at[kMaxEvalAt]; int K; int max_cnt;
  double* dcg;      // [nq][K]
};
__global__ void __launch_bounds__(kMetricThreads) k_ndcg(NdcgArgs a) {
  extern __shared__ __align__(16) unsigned char nd_smem[];
  double* s_score = reinterpret_cast<double*>(nd_smem);
  __shared__ double s_w[8];
  const int q = blockIdx.x;
  const int start = a.qb[q], cnt = a.qb[q + 1] - start;
  for (int d = threadIdx.x; d < cnt; d += blockDim.x) s_score[d] = a.score[start + d];
  __syncthreads();
  double part[kMaxEvalAt];
  #pragma unroll
  for (int j = 0; j < kMaxEvalAt; ++j) part[j] = 0.0;
  const int kmax = a.eval_at[a.K - 1];
  for (int d = threadIdx.x; d < cnt; d += blockDim.x) {
    const double sd = s_score[d];
    int rank = 0;
    for (int e = 0; e < cnt; ++e) {
      const double se = s_score[e];
      rank += (se > sd) || (se == sd && e < d);
    }
    if (rank < kmax) {
      const int lab = static_cast<int>(a.label[start + d]);
      const double g = (lab >= 0 && lab < a.num_gain) ? a.label_gain[lab] : 0.0;
      const double term = g / log2(2.0 + rank);
      #pragma unroll
      for (int j = 0; j < kMaxEvalAt; ++j)
        if (j < a.K && rank < a.eval_at[j]) part[j] += term;
    }
  }
  for (int j = 0; j < a.K; ++j) {
    const double t = BlockSumD(part[j], s_w);
    if (threadIdx.x == 0) a.dcg[static_cast<size_t>(q) * a.K + j] = t;
  }
}
// ---- MAP@k (upstream map_metric.hpp, MapMetric::CalMapAtK).  One CTA per query: documents ranked as by std::stable_sort on falling
// score; every relevant document (label > 0.5) at rank r contributes (relevant documents up to and including r) / (r + 1) to every
// position k > r; the sum is divided by min(relevant documents, k), and a query without relevant documents scores 1.
// out[q * K + j] = AP@eval_at[j].  (NdcgArgs: label_gain / num_gain unused.)
__global__ void __launch_bounds__(kMetricThreads) k_map(NdcgArgs a) {
  extern __shared__ __align__(16) unsigned char nd_smem[];
  double* s_score = reinterpret_cast<double*>(nd_smem);
  int* s_rank = reinterpret_cast<int*>(s_score + a.max_cnt);      // -1 for an irrelevant document
  __shared__ double s_w[8];
  const int q = blockIdx.x;
  const int start = a.qb[q], cnt = a.qb[q + 1] - start;
  for (int d = threadIdx.x; d < cnt; d += blockDim.x) s_score[d] = a.score[start + d];
  __syncthreads();
  double npos_part = 0.0;
  for (int d = threadIdx.x; d < cnt; d += blockDim.x) {
    const bool rel = a.label[start + d] > 0.5f;
    int rank = -1;
    if (rel) {
      const double sd = s_score[d];
      rank = 0;
      for (int e = 0; e < cnt; ++e) {
        const double se = s_score[e];
        rank += (se > sd) || (se == sd && e < d);
      }
      npos_part += 1.0;
    }
    s_rank[d] = rank;
  }
  __syncthreads();
  const double npos = BlockSumD(npos_part, s_w);
  double part[kMaxEvalAt];
  #pragma unroll
  for (int j = 0; j < kMaxEvalAt; ++j) part[j] = 0.0;
  for (int d = threadIdx.x; d < cnt; d += blockDim.x) {
    const int r = s_rank[d];
    if (r < 0) continue;
    int hits = 0;
    for (int e = 0; e < cnt; ++e) { const int re = s_rank[e]; hits += (re >= 0 && re <= r); }
    const double term = static_cast<double>(hits) / (r + 1.0f);
    #pragma unroll
    for (int j = 0; j < kMaxEvalAt; ++j)
      if (j < a.K && r < min(a.eval_at[j], cnt)) part[j] += term;
  }
  for (int j = 0; j < a.K; ++j) {
    const double t = BlockSumD(part[j], s_w);
    if (threadIdx.x == 0) {
      const double cur_k = static_cast<double>(min(a.eval_at[j], cnt));
      a.dcg[static_cast<size_t>(q) * a.K + j] = npos > 0.0 ? t / fmin(npos, cur_k) : 1.0;
    }
  }
}

// out[j] = sum_q w_q * (max_dcg > 0 ? dcg / max_dcg : 1), out[K] = sum_q w_q; one CTA, fixed order
// (inv_max_dcg == nullptr: the per-query values are final already — MAP)
__global__ void __launch_bounds__(kMetricThreads) k_ndcg_final(const double* __restrict__ dcg, const double* __restrict__ inv_max_dcg,
                                                                const double* __restrict__ qweight, int nq, int K, double* __restrict__ out) {
  __shared__ double s_w[8];
  for (int j = 0; j <= K; ++j) {
    double acc = 0.0;
    for (int q = threadIdx.x; q < nq; q += blockDim.x) {
      const double w = qweight ? qweight[q] : 1.0;
      if (j == K) { acc += w; continue; }
      if (inv_max_dcg == nullptr) { acc += w * dcg[static_cast<size_t>(q) * K + j]; continue; }
      const double inv = inv_max_dcg[static_cast<size_t>(q) * K + j];
      acc += w * (inv > 0.0 ? dcg[static_cast<size_t>(q) * K + j] * inv : 1.0);
    }
    const double t = BlockSumD(acc, s_w);
    if (threadIdx.x == 0) out[j] = t;
  }
}

}  // namespace lgb
