// Device kernels of the training hot path (K3-K8).  Algorithm specification: SURVEY.md Appendix A.5-A.9 (upstream
// objective / feature_histogram / data_partition / gbdt); the kernels are designed for sm_100a from that specification.
//
// Conventions
//  * Histograms are exact integers in every mode.  Quantized mode accumulates the int8 (grad, hess) pairs; float mode
//    first converts float32 grad/hess to 19-bit fixed point with power-of-two scales (one scale per iteration from
//    max|g|, max|h|), so sums are order-independent: the result does not depend on the number of GPUs or on atomic
//    ordering, and the int64 global histogram converts back to float exactly (scale is 2^k).
//  * Global histogram entry = (int64 sum_g, int64 sum_h) per (feature, bin), 256 bin slots per feature.
//  * All per-tree control flow lives in device memory (BatchCtrl / NodeDev, batch_tree.cuh) so a whole tree is enqueued
//    with no host round trip; kernels return immediately once ctrl->done is set.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "booster.h"
#include "hist_kernel.cuh"
#include "softmath.cuh"

namespace lgb {

#define KEPS ((double)1e-15f)
#define KMINSCORE (-__longlong_as_double(0x7ff0000000000000LL))
constexpr int kFixBits = 19;                       // float mode: |fixed-point value| <= 2^19 per row
constexpr int kFixRowsPerCopy = 4000;              // 4000 * 2^19 < 2^31
constexpr int kPartBlocks = 592;                   // 4 x 148 SMs
constexpr int kPartThreads = 256;

__device__ __forceinline__ int RoundIntD(double x) { return static_cast<int>(x + 0.5f); }
__device__ __forceinline__ int SignD(double x) { return (x > 0.0) - (x < 0.0); }

// ------------------------------------------------------------------------------------------------ K3 gradients
// Row-wise objectives as device functions: the stand-alone gradient kernels and the fused apply-tree kernel (batch_tree.cuh)
// must produce the same bits.
__device__ __forceinline__ void GradBinaryRow(double score, float label, bool has_w, float w, double sigmoid, double lw_neg,
                                              double lw_pos, float* g, float* h) {
  const int is_pos = label > 0.0f;
  const int lab = is_pos ? 1 : -1;
  const double lw = is_pos ? lw_pos : lw_neg;
  const double response = -lab * sigmoid / (1.0f + SoftExp(lab * sigmoid * score));
  const double abs_response = fabs(response);
  if (!has_w) {
    *g = static_cast<float>(response * lw);
    *h = static_cast<float>(abs_response * (sigmoid - abs_response) * lw);
  } else {
    *g = static_cast<float>(response * lw * w);
    *h = static_cast<float>(abs_response * (sigmoid - abs_response) * lw * w);
  }
}
__device__ __forceinline__ void GradL2Row(double score, float label, bool has_w, float w, float* g, float* h) {
  if (!has_w) {
    *g = static_cast<float>(score - label);
    *h = 1.0f;
  } else {
    *g = static_cast<float>(static_cast<float>(score - label) * w);
    *h = static_cast<float>(w);
  }
}

// pos_class < 0: positive = label > 0 (binary); otherwise positive = "label == pos_class" (one class of multiclassova)
__global__ void k_grad_binary(const double* __restrict__ score, const float* __restrict__ label, const float* __restrict__ weight,
                              float* __restrict__ grad, float* __restrict__ hess, int n, double sigmoid, double lw_neg, double lw_pos,
                              int pos_class = -1) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const float lab = pos_class < 0 ? label[i] : (static_cast<int>(label[i]) == pos_class ? 1.0f : 0.0f);
    GradBinaryRow(score[i], lab, weight != nullptr, weight ? weight[i] : 0.0f, sigmoid, lw_neg, lw_pos, grad + i, hess + i);
  }
}

__global__ void k_grad_l2(const double* __restrict__ score, const float* __restrict__ label, const float* __restrict__ weight,
                          float* __restrict__ grad, float* __restrict__ hess, int n) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    GradL2Row(score[i], label[i], weight != nullptr, weight ? weight[i] : 0.0f, grad + i, hess + i);
}

// Huber / Fair / Poisson / Gamma / Tweedie (upstream regression_objective.hpp); kind = Booster::Obj.  p0: alpha | fair_c |
// exp(poisson_max_delta_step) | - | tweedie_variance_power.
// kinds 10 / 11 / 12: regression_l1 (sign of the error), quantile (p0 = alpha, kept as a float as upstream does), mape (sign of the
// error times label_weight = 1 / max(1, |label|) [* weight]); unit hessians (the row weight when there are weights).
__global__ void k_grad_point(int kind, const double* __restrict__ score, const float* __restrict__ label, const float* __restrict__ weight,
                             float* __restrict__ grad, float* __restrict__ hess, int n, double p0, const float* __restrict__ label_weight) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const double sc = score[i];
    const double y = label[i];
    double g, h;
    if (kind == 15) {           // cross_entropy (upstream xentropy_objective.hpp): labels are probabilities
      const double z = 1.0f / (1.0f + SoftExp(-sc));
      g = z - y;
      h = z * (1.0f - z);
    } else if (kind >= 10) {
      const double diff = sc - y;
      if (kind == 10) {
        g = SignD(diff);
      } else if (kind == 11) {
        const float delta = static_cast<float>(diff);
        const float alpha = static_cast<float>(p0);
        g = delta >= 0 ? (1.0f - alpha) : -alpha;
      } else {
        grad[i] = static_cast<float>(SignD(diff) * label_weight[i]);      // label_weight already carries the row weight
        hess[i] = weight == nullptr ? 1.0f : weight[i];
        continue;
      }
      h = 1.0;
    } else if (kind == 4) {            // huber
      const double diff = sc - y;
      g = fabs(diff) <= p0 ? diff : SignD(diff) * p0;
      h = 1.0;
    } else if (kind == 5) {     // fair
      const double x = sc - y;
      g = p0 * x / (fabs(x) + p0);
      h = p0 * p0 / ((fabs(x) + p0) * (fabs(x) + p0));
    } else if (kind == 6) {     // poisson
      const double e = SoftExp(sc);
      g = e - y;
      h = e * p0;
    } else if (kind == 7) {     // gamma
      const double e = SoftExp(-sc);
      g = 1.0 - y * e;
      h = y * e;
    } else {                    // tweedie
      const double e1 = SoftExp((1 - p0) * sc), e2 = SoftExp((2 - p0) * sc);
      g = -y * e1 + e2;
      h = -y * (1 - p0) * e1 + (2 - p0) * e2;
    }
    if (weight == nullptr) { grad[i] = static_cast<float>(g); hess[i] = static_cast<float>(h); }
    else { grad[i] = static_cast<float>(g * weight[i]); hess[i] = static_cast<float>(h * weight[i]); }
  }
}

// Multiclass softmax (upstream multiclass_objective.hpp, MulticlassSoftmax::GetGradients).  score / grad / hess are class-major
// [K][n].  Per row: Common::Softmax over its K scores (subtract the maximum, exp, divide by the sum; exp is the software
// function the oracle shares), g_k = p_k - [label == k], h_k = K / (K - 1) * p_k * (1 - p_k), times the row weight.
constexpr int kMaxClasses = 64;
__global__ void k_grad_multiclass(const double* __restrict__ score, const float* __restrict__ label, const float* __restrict__ weight,
                                  float* __restrict__ grad, float* __restrict__ hess, int n, int K, double factor) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    double rec[kMaxClasses];
    double wmax = score[i];
    for (int k = 0; k < K; ++k) { rec[k] = score[static_cast<size_t>(k) * n + i]; wmax = fmax(rec[k], wmax); }
    double wsum = 0.0;
    for (int k = 0; k < K; ++k) { rec[k] = SoftExp(rec[k] - wmax); wsum += rec[k]; }
    const int lab = static_cast<int>(label[i]);
    const bool has_w = weight != nullptr;
    const float w = has_w ? weight[i] : 0.0f;
    for (int k = 0; k < K; ++k) {
      const double p = rec[k] / wsum;
      const size_t idx = static_cast<size_t>(k) * n + i;
      if (!has_w) {
        grad[idx] = lab == k ? static_cast<float>(p - 1.0f) : static_cast<float>(p);
        hess[idx] = static_cast<float>(factor * p * (1.0f - p));
      } else {
        grad[idx] = lab == k ? static_cast<float>((p - 1.0f) * w) : static_cast<float>(p * w);
        hess[idx] = static_cast<float>(factor * p * (1.0f - p) * w);
      }
    }
  }
}

// LambdaRank-NDCG.  One CTA per query.  Step 1: stable rank of every document by score (descending) by counting, which
// equals std::stable_sort's order.  Step 2 (k_grad_lambdarank): the pair loop, with every floating-point sum taken in the
// sequential loop's order so float32 lambdas / hessians and the double sum behind the log2 normalisation are bit-identical.
struct LambdarankArgs {
  const double* score; const float* label; const float* weight; const int* qb;
  const double* inv_max_dcg; const double* label_gain; const double* discount; const double* sigmoid_table;
  float* grad; float* hess;
  double sigmoid, min_sigmoid_input, max_sigmoid_input, idx_factor;
  int truncation_level, norm, table_size, max_cnt;
  const int* qlist;      // queries of this launch (nullptr: query = blockIdx.x); launches are split by query size so that short
                         // queries do not pay for the shared memory of the longest one
};

__device__ __forceinline__ double LrSigmoid(const LambdarankArgs& a, double s) {
  if (s <= a.min_sigmoid_input) return a.sigmoid_table[0];
  if (s >= a.max_sigmoid_input) return a.sigmoid_table[a.table_size - 1];
  return a.sigmoid_table[static_cast<size_t>((s - a.min_sigmoid_input) * a.idx_factor)];
}

// pair (i < j in sorted order): computes p_lambda, p_hessian; returns false if the pair is skipped
__device__ __forceinline__ bool LrPair(const LambdarankArgs& a, const double* s_score, const float* s_label, int i, int j,
                                       double inverse_max_dcg, bool norm_on, double* p_lambda, double* p_hessian, bool* i_is_high) {
  if (s_label[i] == s_label[j]) return false;
  int high_rank, low_rank;
  if (s_label[i] > s_label[j]) { high_rank = i; low_rank = j; } else { high_rank = j; low_rank = i; }
  const double high_score = s_score[high_rank], low_score = s_score[low_rank];
  const double high_label_gain = a.label_gain[static_cast<int>(s_label[high_rank])];
  const double low_label_gain = a.label_gain[static_cast<int>(s_label[low_rank])];
  const double high_discount = a.discount[high_rank], low_discount = a.discount[low_rank];
  const double delta_score = high_score - low_score;
  const double dcg_gap = high_label_gain - low_label_gain;
  const double paired_discount = fabs(high_discount - low_discount);
  double delta_pair_NDCG = dcg_gap * paired_discount * inverse_max_dcg;
  if (norm_on) delta_pair_NDCG /= (0.01f + fabs(delta_score));
  double pl = LrSigmoid(a, delta_score);
  double ph = pl * (1.0f - pl);
  pl *= -a.sigmoid * delta_pair_NDCG;
  ph *= a.sigmoid * a.sigmoid * delta_pair_NDCG;
  *p_lambda = pl; *p_hessian = ph; *i_is_high = (high_rank == i);
  return true;
}

// Every pair (i, j), i < min(cnt - 1, truncation), j > i, is evaluated ONCE, by the thread that owns document j in the
// iteration of the (sequential) outer loop over i; the three order-sensitive accumulations of the CPU loop are kept in its
// order: document j's float sums take their (i, j) terms in increasing i (own registers-in-shared), document i's float sums
// take the (i, j) terms in increasing j (one thread walks the row that was just staged in shared memory), and the double
// sum of lambdas for the normalisation walks the same row on another thread.
__global__ void __launch_bounds__(256) k_grad_lambdarank(LambdarankArgs a) {
  extern __shared__ __align__(16) unsigned char lr_smem[];
  const int q = a.qlist ? a.qlist[blockIdx.x] : static_cast<int>(blockIdx.x);
  const int start = a.qb[q];
  const int cnt = a.qb[q + 1] - start;
  double* s_score = reinterpret_cast<double*>(lr_smem);                 // [max_cnt] sorted by rank
  double* s_2pl = s_score + a.max_cnt;                                  // staged row: 2 * pair lambda (0 for pairs that do not count)
  float* s_label = reinterpret_cast<float*>(s_2pl + a.max_cnt);        // [max_cnt]
  float* s_lam = s_label + a.max_cnt;
  float* s_hes = s_lam + a.max_cnt;
  float* s_di = s_hes + a.max_cnt;                                      // staged row: the pair's signed lambda as document i sees it
  float* s_hf = s_di + a.max_cnt;                                       // staged row: the pair's hessian
  int* s_doc = reinterpret_cast<int*>(s_hf + a.max_cnt);               // rank -> doc
  __shared__ double s_sum;
  const double* sc = a.score + start;
  const float* lb = a.label + start;
  for (int d = threadIdx.x; d < cnt; d += blockDim.x) {
    const double sd = sc[d];
    int rank = 0;
    for (int e = 0; e < cnt; ++e) {
      const double se = sc[e];
      rank += (se > sd) || (se == sd && e < d);
    }
    s_score[rank] = sd; s_label[rank] = lb[d]; s_doc[rank] = d;
    s_lam[rank] = 0.0f; s_hes[rank] = 0.0f;
  }
  if (threadIdx.x == 0) s_sum = 0.0;
  __syncthreads();
  const double inverse_max_dcg = a.inv_max_dcg[q];
  const double best_score = s_score[0];
  int worst_idx = cnt - 1;
  if (worst_idx > 0 && s_score[worst_idx] == KMINSCORE) worst_idx -= 1;
  const double worst_score = s_score[worst_idx];
  const bool norm_on = a.norm && best_score != worst_score;
  const int T = a.truncation_level;
  const int imax = min(cnt - 1, T);      // i ranges over [0, imax)
  for (int i = 0; i < imax; ++i) {
    if (s_score[i] == KMINSCORE) continue;
    for (int j = i + 1 + threadIdx.x; j < cnt; j += blockDim.x) {
      float di = 0.0f, hf = 0.0f;
      double two_pl = 0.0;
      if (s_score[j] != KMINSCORE) {
        double pl, ph; bool i_high;
        if (LrPair(a, s_score, s_label, i, j, inverse_max_dcg, norm_on, &pl, &ph, &i_high)) {
          const float plf = static_cast<float>(pl);
          hf = static_cast<float>(ph);
          two_pl = 2 * pl;
          di = i_high ? plf : -plf;
          // document j is the lower one when i is the higher one
          s_lam[j] = i_high ? s_lam[j] - plf : s_lam[j] + plf;
          s_hes[j] += hf;
        }
      }
      // pairs that do not count are staged as zeros: x + 0 == x, so the walks below need no branch and pipeline
      s_di[j] = di; s_hf[j] = hf; s_2pl[j] = two_pl;
    }
    __syncthreads();
    if (threadIdx.x == (i & 31)) {               // warp 0: document i's sums over its row, in increasing j
      float lam = s_lam[i], hes = s_hes[i];
      for (int j = i + 1; j < cnt; ++j) { lam += s_di[j]; hes += s_hf[j]; }
      s_lam[i] = lam; s_hes[i] = hes;
    } else if (a.norm && threadIdx.x == 32 + (i & 31)) {   // warp 1: the double sum for the normalisation, same order
      double sum = s_sum;
      for (int j = i + 1; j < cnt; ++j) sum -= s_2pl[j];
      s_sum = sum;
    }
    __syncthreads();
  }
  double nf = -1.0;
  if (a.norm) {
    const double sum_lambdas = s_sum;
    nf = (sum_lambdas > 0) ? SoftLog2(1 + sum_lambdas) / sum_lambdas : -1.0;
  }
  for (int r = threadIdx.x; r < cnt; r += blockDim.x) {
    float lam = s_lam[r], hes = s_hes[r];
    if (nf >= 0.0) {
      lam = static_cast<float>(lam * nf);
      hes = static_cast<float>(hes * nf);
    }
    const int d = start + s_doc[r];
    if (a.weight != nullptr) {
      lam = static_cast<float>(lam * a.weight[d]);
      hes = static_cast<float>(hes * a.weight[d]);
    }
    a.grad[d] = lam;
    a.hess[d] = hes;
  }
}

// rank_xendcg (upstream RankXENDCG::GetGradientsForOneQuery): one CTA per query.  The ground-truth distribution is 2^label minus a
// uniform draw from the QUERY's own LCG stream (one draw per document per call, in document order), the model's is the softmax of the
// scores; gradient = the first three terms of the expansion, hessian = rho (1 - rho).  The element-wise parts run on all threads;
// the draws and the four sums (softmax denominator, ground-truth denominator, first- and second-order sums) run on thread 0 in
// document order, as the CPU loop does, so float32 results are bit-identical.  Shared: 3 doubles per document.
struct XendcgArgs {
  const double* score; const float* label; const float* weight; const int* qb; unsigned int* rand_state; float* grad; float* hess; int max_cnt;
};
__global__ void __launch_bounds__(256) k_grad_xendcg(XendcgArgs a) {
  extern __shared__ __align__(16) unsigned char xe_smem[];
  double* s_rho = reinterpret_cast<double*>(xe_smem);
  double* s_par = s_rho + a.max_cnt;
  __shared__ double s_red[8];
  __shared__ double s_val;
  const int q = blockIdx.x;
  const int start = a.qb[q], cnt = a.qb[q + 1] - start;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (cnt <= 1) {
    if (tid < cnt) { a.grad[start + tid] = 0.0f; a.hess[start + tid] = 0.0f; }
    return;
  }
  // softmax: maximum (any order), exp, then the sum in document order
  double wmax = -INFINITY;
  for (int i = tid; i < cnt; i += blockDim.x) wmax = fmax(wmax, a.score[start + i]);
  for (int o = 16; o > 0; o >>= 1) wmax = fmax(wmax, __shfl_xor_sync(0xffffffffu, wmax, o));
  if (lane == 0) s_red[warp] = wmax;
  __syncthreads();
  wmax = s_red[0];
  #pragma unroll
  for (int w = 1; w < 8; ++w) wmax = fmax(wmax, s_red[w]);
  for (int i = tid; i < cnt; i += blockDim.x) s_rho[i] = SoftExp(a.score[start + i] - wmax);
  __syncthreads();
  if (tid == 0) {
    double wsum = 0.0;
    for (int i = 0; i < cnt; ++i) wsum += s_rho[i];
    s_val = wsum;
    // Phi(label, draw) and the ground-truth denominator
    unsigned int x = a.rand_state[q];
    double den = 0.0;
    for (int i = 0; i < cnt; ++i) {
      x = 214013u * x + 2531011u;
      const float u = static_cast<float>((x >> 16) & 0x7FFF) / 32768.0f;
      const double p = static_cast<double>(1 << static_cast<int>(a.label[start + i])) - u;
      s_par[i] = p;
      den += p;
    }
    a.rand_state[q] = x;
    s_red[0] = 1. / fmax(KEPS, den);
  }
  __syncthreads();
  const double wsum = s_val, inv_den = s_red[0];
  __syncthreads();
  float lam[8];      // this thread's documents tid, tid + 256, ... (queries of up to 2048 documents keep them in registers)
  int nd = 0;
  for (int i = tid; i < cnt; i += blockDim.x, ++nd) {
    const double rho = s_rho[i] / wsum;
    s_rho[i] = rho;
    const double term = -s_par[i] * inv_den + rho;
    if (nd < 8) lam[nd] = static_cast<float>(term); else a.grad[start + i] = static_cast<float>(term);
    s_par[i] = term / (1. - rho);
  }
  __syncthreads();
  if (tid == 0) { double s1 = 0.0; for (int i = 0; i < cnt; ++i) s1 += s_par[i]; s_val = s1; }
  __syncthreads();
  const double sum_l1 = s_val;
  __syncthreads();
  nd = 0;
  for (int i = tid; i < cnt; i += blockDim.x, ++nd) {
    const double term = s_rho[i] * (sum_l1 - s_par[i]);
    if (nd < 8) lam[nd] += static_cast<float>(term); else a.grad[start + i] += static_cast<float>(term);
    s_par[i] = term / (1. - s_rho[i]);
  }
  __syncthreads();
  if (tid == 0) { double s2 = 0.0; for (int i = 0; i < cnt; ++i) s2 += s_par[i]; s_val = s2; }
  __syncthreads();
  const double sum_l2 = s_val;
  nd = 0;
  for (int i = tid; i < cnt; i += blockDim.x, ++nd) {
    const double rho = s_rho[i];
    float g = nd < 8 ? lam[nd] : a.grad[start + i];
    g += static_cast<float>(rho * (sum_l2 - s_par[i]));
    float h = static_cast<float>(rho * (1.0 - rho));
    if (a.weight != nullptr) { g = static_cast<float>(g * a.weight[start + i]); h = static_cast<float>(h * a.weight[start + i]); }
    a.grad[start + i] = g; a.hess[start + i] = h;
  }
}

__global__ void k_add_constant(double* score, int n, double v) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) score[i] += v;
}

// max|g|, max|h| over rows (non-negative float bit patterns order like unsigned ints)
template <typename Ctrl>
__global__ void k_absmax(const float* __restrict__ grad, const float* __restrict__ hess, int n, Ctrl* ctrl) {
  float mg = 0.0f, mh = 0.0f;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    mg = fmaxf(mg, fabsf(grad[i]));
    mh = fmaxf(mh, fabsf(hess[i]));
  }
  for (int o = 16; o > 0; o >>= 1) {
    mg = fmaxf(mg, __shfl_xor_sync(0xffffffffu, mg, o));
    mh = fmaxf(mh, __shfl_xor_sync(0xffffffffu, mh, o));
  }
  if ((threadIdx.x & 31) == 0) {
    atomicMax(&ctrl->max_abs_g_bits, __float_as_uint(mg));
    atomicMax(&ctrl->max_abs_h_bits, __float_as_uint(mh));
  }
}

template <typename Ctrl>
__global__ void k_scales(Ctrl* ctrl, LearnerParams lp) {
  const double max_g = __uint_as_float(ctrl->max_abs_g_bits);
  const double max_h = __uint_as_float(ctrl->max_abs_h_bits);
  if (lp.quantized) {
    ctrl->grad_scale = max_g / static_cast<double>(lp.num_grad_quant_bins / 2);
    ctrl->hess_scale = lp.const_hessian ? max_h : max_h / static_cast<double>(lp.num_grad_quant_bins);
    ctrl->inv_grad_scale = 1.0f / ctrl->grad_scale;
    ctrl->inv_hess_scale = 1.0f / ctrl->hess_scale;
  } else {
    // power-of-two fixed-point scales: |x| < 2^e  =>  |x * 2^(kFixBits - e)| < 2^kFixBits
    int eg = 0, eh = 0;
    if (max_g > 0.0) frexp(max_g, &eg);
    if (max_h > 0.0) frexp(max_h, &eh);
    ctrl->inv_grad_scale = ldexp(1.0, kFixBits - eg);
    ctrl->inv_hess_scale = ldexp(1.0, kFixBits - eh);
    ctrl->grad_scale = ldexp(1.0, eg - kFixBits);
    ctrl->hess_scale = ldexp(1.0, eh - kFixBits);
  }
}

template <typename Ctrl>
__global__ void k_quantize(const float* __restrict__ grad, const float* __restrict__ hess, int n, const Ctrl* __restrict__ ctrl,
                           LearnerParams lp, unsigned int iter, unsigned int row_offset, uint16_t* __restrict__ gh16,
                           int2* __restrict__ ghfix, const unsigned int* __restrict__ bag_bits) {
  const double inv_g = ctrl->inv_grad_scale, inv_h = ctrl->inv_hess_scale;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    if (bag_bits != nullptr && !((bag_bits[i >> 5] >> (i & 31)) & 1u)) {      // outside the bag: adds nothing to any histogram
      if (lp.quantized) gh16[i] = 0; else ghfix[i] = make_int2(0, 0);
      continue;
    }
    const float g = grad[i], h = hess[i];
    if (lp.quantized) {
      signed char qg, qh;
      if (lp.stochastic_rounding) {
        const float rg = HashUniform(lp.seed, iter, row_offset + i, 0);
        const float rh = HashUniform(lp.seed, iter, row_offset + i, 1);
        qg = g >= 0.0f ? static_cast<signed char>(g * inv_g + rg) : static_cast<signed char>(g * inv_g - rg);
        qh = static_cast<signed char>(h * inv_h + rh);
      } else {
        qg = g >= 0.0f ? static_cast<signed char>(g * inv_g + 0.5f) : static_cast<signed char>(g * inv_g - 0.5f);
        qh = static_cast<signed char>(h * inv_h + 0.5f);
      }
      gh16[i] = static_cast<uint16_t>((static_cast<unsigned int>(static_cast<unsigned char>(qg)) << 8) | static_cast<unsigned char>(qh));
    } else {
      ghfix[i] = make_int2(__double2int_rn(g * inv_g), __double2int_rn(h * inv_h));
    }
  }
}

// Bagging (upstream src/boosting/bagging.hpp, BaggingHelper): every row draws once from the LCG stream of its 1024-row block
// (seeded bagging_seed + block; the streams run on from one re-bagging to the next) and is in the bag when the 15-bit draw / 32768
// is below bagging_fraction.  One thread per block; writes the bag as a bit per row and adds the bag size to *count.
// label != nullptr: balanced bagging (BalancedBaggingHelper) — positives (label > 0) are kept with pos_fraction, the others with
// `fraction` (= neg_bagging_fraction).
__global__ void k_bag_sample(unsigned int* __restrict__ rand_state, int n, double fraction, unsigned int* __restrict__ bits,
                             int* __restrict__ count, const float* __restrict__ label = nullptr, double pos_fraction = 1.0) {
  const int blk = blockIdx.x * blockDim.x + threadIdx.x;
  if (blk >= (n + 1023) / 1024) return;
  unsigned int x = rand_state[blk];
  int cnt = 0;
  for (int w = 0; w < 32; ++w) {
    unsigned int word = 0;
    for (int j = 0; j < 32; ++j) {
      if (blk * 1024 + w * 32 + j < n) {
        x = 214013u * x + 2531011u;
        const float u = static_cast<float>((x >> 16) & 0x7FFF) / 32768.0f;
        const double fr = (label != nullptr && label[blk * 1024 + w * 32 + j] > 0.0f) ? pos_fraction : fraction;
        if (static_cast<double>(u) < fr) { word |= 1u << j; ++cnt; }
      }
    }
    bits[blk * 32 + w] = word;
  }
  rand_state[blk] = x;
  atomicAdd(count, cnt);
}

// ------------------------------------------------------------------------------------------------ K5 + K6 split search
__device__ __forceinline__ double ThresholdL1D(double s, double l1) {
  double reg = fabs(s) - l1;
  if (reg < 0.0) reg = 0.0;
  return SignD(s) * reg;
}
__device__ __forceinline__ double LeafOutputD(double g, double h, const LearnerParams& p) {
  double ret = (p.l1 > 0.0) ? -ThresholdL1D(g, p.l1) / (h + p.l2) : -g / (h + p.l2);
  if (p.max_delta_step > 0.0 && fabs(ret) > p.max_delta_step) ret = SignD(ret) * p.max_delta_step;
  return ret;
}
__device__ __forceinline__ double LeafGainD(double g, double h, const LearnerParams& p) {
  if (p.max_delta_step <= 0.0) {
    const double sg = (p.l1 > 0.0) ? ThresholdL1D(g, p.l1) : g;
    return (sg * sg) / (h + p.l2);
  }
  const double out = LeafOutputD(g, h, p);
  const double sg = (p.l1 > 0.0) ? ThresholdL1D(g, p.l1) : g;
  return -(2.0 * sg * out + (h + p.l2) * out * out);
}
__device__ __forceinline__ double SplitGainD(double lg, double lh, double rg, double rh, const LearnerParams& p) {
  return LeafGainD(lg, lh, p) + LeafGainD(rg, rh, p);
}

// Monotone constraints, method "basic" (upstream monotone_constraints.hpp BasicLeafConstraints, feature_histogram.hpp USE_MC): a leaf
// carries an output range [cmin, cmax]; while any column is constrained, every candidate's gain is taken at the children's outputs
// clamped to that range (GetLeafGainGivenOutput), a split that breaks the searched feature's direction has gain 0, and the outputs
// handed on are the clamped ones.
struct McCtx { int on; int mono; double cmin, cmax; };
__device__ __forceinline__ double LeafGainGivenOutputD(double g, double h, const LearnerParams& p, double out) {
  const double sg = (p.l1 > 0.0) ? ThresholdL1D(g, p.l1) : g;
  return -(2.0 * sg * out + (h + p.l2) * out * out);
}
__device__ __forceinline__ double LeafOutputX(double g, double h, const LearnerParams& p, const McCtx& mc) {
  double ret = LeafOutputD(g, h, p);
  if (mc.on) { if (ret < mc.cmin) ret = mc.cmin; else if (ret > mc.cmax) ret = mc.cmax; }
  return ret;
}
__device__ __forceinline__ double SplitGainX(double lg, double lh, double rg, double rh, const LearnerParams& p, const McCtx& mc) {
  if (!mc.on) return LeafGainD(lg, lh, p) + LeafGainD(rg, rh, p);
  const double lo = LeafOutputX(lg, lh, p, mc), ro = LeafOutputX(rg, rh, p, mc);
  if ((mc.mono > 0 && lo > ro) || (mc.mono < 0 && lo < ro)) return 0.0;
  return LeafGainGivenOutputD(lg, lh, p, lo) + LeafGainGivenOutputD(rg, rh, p, ro);
}
// LeafConstraintsBase::ComputeMonotoneSplitGainPenalty: the factor on the gain of a split on a constrained feature at this depth
__device__ __forceinline__ double MonotonePenaltyD(int depth, double penalization) {
  if (penalization >= depth + 1.) return KEPS;
  if (penalization <= 1.) return 1. - penalization / pow(2., static_cast<double>(depth)) + KEPS;
  return 1. - pow(2., penalization - 1. - depth) + KEPS;
}

__device__ __forceinline__ bool SplitBetter(double ga, int fa, double gb, int fb) {
  // upstream SplitInfo::operator>: larger gain; equal gain -> smaller real feature index (-1 = +inf)
  const int a = fa < 0 ? 0x7fffffff : fa, b = fb < 0 ? 0x7fffffff : fb;
  if (ga != gb) return ga > gb;
  return a < b;
}

__device__ __forceinline__ SplitDev EmptySplit(int f, int real) {
  SplitDev s;
  s.gain = KMINSCORE; s.feature = f; s.real_feature = real; s.threshold = 0; s.default_left = 1;
  s.left_count = s.right_count = 0; s.left_output = s.right_output = 0.0;
  s.left_sum_g = s.left_sum_h = s.right_sum_g = s.right_sum_h = 0.0;
  s.left_int_g = s.left_int_h = s.right_int_g = s.right_int_h = 0;
  s.missing_type = 0; s.nan_bin = -1;
  s.is_cat = 0; s.monotone_type = 0;
  #pragma unroll
  for (int i = 0; i < 8; ++i) s.cat_bits[i] = 0u;
  return s;
}

// 256-thread sub-blocks of a CTA synchronise on named barriers instead of __syncthreads() (the partition kernel's tiles, the replay).
constexpr int kFindThreads = 256;
constexpr int kMaxFindSubs = 4;
__device__ __forceinline__ void SubSync(int sub) { asm volatile("bar.sync %0, %1;" ::"r"(sub + 1), "n"(kFindThreads) : "memory"); }

struct Cand { double gain; int b; long long sg, sh; int cnt; };   // sums of the scanned side (right: reverse, left: forward)

// block-wide argmax of candidates; ties on gain go to the larger (reverse scan) or smaller (forward scan) bin, which is
// what a sequential scan with a strict '>' keeps
__device__ __forceinline__ Cand BlockBestCand(Cand c, bool prefer_larger_b, Cand* s_c /*[8]*/, int sub) {
  const int lane = threadIdx.x & 31, warp = (threadIdx.x & (kFindThreads - 1)) >> 5;
  auto better = [&](const Cand& t, const Cand& cur) {
    return (t.gain > cur.gain) || (t.gain == cur.gain && t.b >= 0 && (cur.b < 0 || (prefer_larger_b ? t.b > cur.b : t.b < cur.b)));
  };
  for (int o = 16; o > 0; o >>= 1) {
    Cand t;
    t.gain = __shfl_xor_sync(0xffffffffu, c.gain, o);
    t.b = __shfl_xor_sync(0xffffffffu, c.b, o);
    t.sg = __shfl_xor_sync(0xffffffffu, c.sg, o);
    t.sh = __shfl_xor_sync(0xffffffffu, c.sh, o);
    t.cnt = __shfl_xor_sync(0xffffffffu, c.cnt, o);
    if (better(t, c)) c = t;
  }
  SubSync(sub);
  if (lane == 0) s_c[warp] = c;
  SubSync(sub);
  Cand r = s_c[0];
  #pragma unroll
  for (int w = 1; w < 8; ++w) if (better(s_c[w], r)) r = s_c[w];
  return r;
}

// One 256-thread CTA searches one feature of one leaf, one bin per thread.  Implements FindBestThresholdSequentially /
// ...Int (reverse scan, plus a forward scan when the feature has a NaN bin and more than two bins) with block-wide prefix
// sums: the sequential `continue` / `break` conditions are monotone in the scan position, so they become per-threshold
// predicates.  g_in/h_in: this thread's bin of the leaf's histogram.  Result valid in thread 0.
template <bool QUANT, bool MC = false>
__device__ void BlockFindBest(long long g_in, long long h_in, const FeatureMetaDev fm, int f, bool root, long long tot_g,
                              long long tot_h, double leaf_sum_g, double leaf_sum_h, int num_data, const LearnerParams& P,
                              double gscale, double hscale, SplitDev* out, unsigned char* splittable_out,
                              long long* out_tot_g, long long* out_tot_h, int sub, const McCtx mc = McCtx{0, 0, 0.0, 0.0}, int depth = 0) {
  __shared__ long long s_g_all[kMaxFindSubs][8], s_h_all[kMaxFindSubs][8];
  __shared__ int s_c_all[kMaxFindSubs][8];
  __shared__ Cand s_cand_all[kMaxFindSubs][8];
  __shared__ int s_flag_all[kMaxFindSubs];
  long long* s_g = s_g_all[sub]; long long* s_h = s_h_all[sub]; int* s_c = s_c_all[sub]; Cand* s_cand = s_cand_all[sub];
  int& s_flag = s_flag_all[sub];
  const int b = threadIdx.x & (kFindThreads - 1), lane = b & 31, warp = b >> 5;
  const int num_bin = fm.num_bin;
  // two scans when the feature has a missing type and more than two bins.  NaN type: the last bin is left out of both scans
  // (NA_AS_MISSING); Zero type (zero_as_missing): the default bin is neither accumulated nor a threshold (SKIP_DEFAULT_BIN), and the
  // forward scan starts at bin 1 when the most frequent bin is bin 0 (upstream keeps that bin out of its histogram)
  const bool zero_type = fm.missing_type == kMissingZero;
  const bool two_scans = (num_bin > 2 && (fm.missing_type == kMissingNaN || zero_type));
  const int nan_bin = !two_scans ? -1 : zero_type ? fm.default_bin : num_bin - 1;      // the bin both scans leave out
  const int fstart = (two_scans && zero_type && fm.most_freq_bin == 0) ? 1 : 0;
  if (b == 0) s_flag = 0;
  // pass 0 (root only): leaf totals = sum of all bins of this feature
  if (root) {
    long long tg = g_in, th = h_in;
    for (int o = 16; o > 0; o >>= 1) { tg += __shfl_xor_sync(0xffffffffu, tg, o); th += __shfl_xor_sync(0xffffffffu, th, o); }
    if (lane == 0) { s_g[warp] = tg; s_h[warp] = th; }
    SubSync(sub);
    tot_g = 0; tot_h = 0;
    #pragma unroll
    for (int w = 0; w < 8; ++w) { tot_g += s_g[w]; tot_h += s_h[w]; }
    leaf_sum_g = static_cast<double>(tot_g) * gscale;
    leaf_sum_h = static_cast<double>(tot_h) * hscale;
    SubSync(sub);
  }
  *out_tot_g = tot_g; *out_tot_h = tot_h;
  double sum_gradient, sum_hessian, cnt_factor;
  if (QUANT) {
    sum_gradient = static_cast<double>(tot_g) * gscale;
    sum_hessian = static_cast<double>(tot_h) * hscale;
    cnt_factor = static_cast<double>(num_data) / static_cast<double>(tot_h);
  } else {
    sum_gradient = leaf_sum_g;
    sum_hessian = leaf_sum_h + 2 * KEPS;
    cnt_factor = num_data / sum_hessian;
  }
  const bool use = b < num_bin && b != nan_bin;
  const long long gv = use ? g_in : 0, hv = use ? h_in : 0;
  const int cv = (!QUANT && use) ? RoundIntD(static_cast<double>(hv) * hscale * cnt_factor) : 0;
  // inclusive block scan
  long long pg = gv, ph = hv; int pc = cv;
  #pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const long long tg = __shfl_up_sync(0xffffffffu, pg, o);
    const long long th = __shfl_up_sync(0xffffffffu, ph, o);
    const int tc = __shfl_up_sync(0xffffffffu, pc, o);
    if (lane >= o) { pg += tg; ph += th; pc += tc; }
  }
  __shared__ long long s_g0_all[kMaxFindSubs], s_h0_all[kMaxFindSubs];
  __shared__ int s_c0_all[kMaxFindSubs];
  if (b == 0) { s_g0_all[sub] = gv; s_h0_all[sub] = hv; s_c0_all[sub] = cv; }      // bin 0, which a forward scan from bin 1 leaves out
  if (lane == 31) { s_g[warp] = pg; s_h[warp] = ph; s_c[warp] = pc; }
  SubSync(sub);
  long long Tg = 0, Th = 0; int Tc = 0;
  #pragma unroll
  for (int w = 0; w < 8; ++w) {
    if (w < warp) { pg += s_g[w]; ph += s_h[w]; pc += s_c[w]; }
    Tg += s_g[w]; Th += s_h[w]; Tc += s_c[w];
  }
  const double gain_shift = LeafGainD(sum_gradient, sum_hessian, P);
  const double min_gain_shift = gain_shift + P.min_gain_to_split;
  bool splittable = false;
  // ---- reverse scan candidate: right = bins [b, top], threshold = b - 1
  const int top = num_bin - 1 - ((two_scans && !zero_type) ? 1 : 0);
  Cand cr; cr.gain = KMINSCORE; cr.b = -1; cr.sg = 0; cr.sh = 0; cr.cnt = 0;
  if (b >= 1 && b <= top && b != nan_bin) {
    const long long rgi = Tg - (pg - gv), rhi = Th - (ph - hv);
    double gain = 0.0; int rc; bool ok = true;
    if (QUANT) {
      rc = RoundIntD(static_cast<double>(rhi) * cnt_factor);
      const double srh = static_cast<double>(rhi) * hscale;
      const int lc = num_data - rc;
      const long long lgi = tot_g - rgi, lhi = tot_h - rhi;
      const double slh = static_cast<double>(lhi) * hscale;
      ok = !(rc < P.min_data_in_leaf || srh < P.min_sum_hessian) && !(lc < P.min_data_in_leaf) && !(slh < P.min_sum_hessian);
      if (ok) gain = (MC ? SplitGainX(static_cast<double>(lgi) * gscale, slh, static_cast<double>(rgi) * gscale, srh, P, mc) : SplitGainD(static_cast<double>(lgi) * gscale, slh, static_cast<double>(rgi) * gscale, srh, P));
    } else {
      rc = Tc - (pc - cv);
      const double rh = KEPS + static_cast<double>(rhi) * hscale;
      const int lc = num_data - rc;
      const double lh = sum_hessian - rh;
      ok = !(rc < P.min_data_in_leaf || rh < P.min_sum_hessian) && !(lc < P.min_data_in_leaf) && !(lh < P.min_sum_hessian);
      if (ok) { const double rg = static_cast<double>(rgi) * gscale; gain = (MC ? SplitGainX(sum_gradient - rg, lh, rg, rh, P, mc) : SplitGainD(sum_gradient - rg, lh, rg, rh, P)); }
    }
    if (ok && gain > min_gain_shift) { splittable = true; cr.gain = gain; cr.b = b; cr.sg = rgi; cr.sh = rhi; cr.cnt = rc; }
  }
  // ---- forward scan candidate (NaN as missing): left = bins [0, b], threshold = b
  Cand cf; cf.gain = KMINSCORE; cf.b = -1; cf.sg = 0; cf.sh = 0; cf.cnt = 0;
  if (two_scans && b >= fstart && b <= num_bin - 2 && b != nan_bin) {
    const long long lgi = pg - (fstart ? s_g0_all[sub] : 0), lhi = ph - (fstart ? s_h0_all[sub] : 0);
    double gain = 0.0; int lc; bool ok = true;
    if (QUANT) {
      lc = RoundIntD(static_cast<double>(lhi) * cnt_factor);
      const double slh = static_cast<double>(lhi) * hscale;
      const int rc = num_data - lc;
      const long long rgi = tot_g - lgi, rhi = tot_h - lhi;
      const double srh = static_cast<double>(rhi) * hscale;
      ok = !(lc < P.min_data_in_leaf || slh < P.min_sum_hessian) && !(rc < P.min_data_in_leaf) && !(srh < P.min_sum_hessian);
      if (ok) gain = (MC ? SplitGainX(static_cast<double>(lgi) * gscale, slh, static_cast<double>(rgi) * gscale, srh, P, mc) : SplitGainD(static_cast<double>(lgi) * gscale, slh, static_cast<double>(rgi) * gscale, srh, P));
    } else {
      lc = pc - (fstart ? s_c0_all[sub] : 0);
      const double lh = KEPS + static_cast<double>(lhi) * hscale;
      const int rc = num_data - lc;
      const double rh = sum_hessian - lh;
      ok = !(lc < P.min_data_in_leaf || lh < P.min_sum_hessian) && !(rc < P.min_data_in_leaf) && !(rh < P.min_sum_hessian);
      if (ok) { const double lg = static_cast<double>(lgi) * gscale; gain = (MC ? SplitGainX(lg, lh, sum_gradient - lg, rh, P, mc) : SplitGainD(lg, lh, sum_gradient - lg, rh, P)); }
    }
    if (ok && gain > min_gain_shift) { splittable = true; cf.gain = gain; cf.b = b; cf.sg = lgi; cf.sh = lhi; cf.cnt = lc; }
  }
  if (splittable) s_flag = 1;
  const Cand best_r = BlockBestCand(cr, true, s_cand, sub);
  Cand best_f = cf;
  if (two_scans) best_f = BlockBestCand(cf, false, s_cand, sub);
  SubSync(sub);
  if (b == 0) {
    const bool any = s_flag != 0;
    const bool accept = !(MC && mc.cmin > mc.cmax);      // an empty output range accepts no candidate (the feature stays "splittable")
    SplitDev s = EmptySplit(f, fm.real_index);
    s.missing_type = fm.missing_type;
    s.nan_bin = (fm.missing_type == kMissingNaN) ? fm.num_bin - 1 : zero_type ? fm.default_bin : -1;      // rows of this bin follow default_left
    for (int pass = 0; pass < 2; ++pass) {
      const Cand& c = pass == 0 ? best_r : best_f;
      if (pass == 1 && !two_scans) break;
      if (!(any && accept && c.b >= 0 && c.gain > s.gain + min_gain_shift)) continue;
      const bool reverse = pass == 0;
      if (QUANT) {
        const long long lgi = reverse ? tot_g - c.sg : c.sg, lhi = reverse ? tot_h - c.sh : c.sh;
        const long long rgi = tot_g - lgi, rhi = tot_h - lhi;
        const double blg = static_cast<double>(lgi) * gscale, blh = static_cast<double>(lhi) * hscale;
        const double brg = static_cast<double>(rgi) * gscale, brh = static_cast<double>(rhi) * hscale;
        s.threshold = reverse ? c.b - 1 : c.b;
        s.left_output = (MC ? LeafOutputX(blg, blh, P, mc) : LeafOutputD(blg, blh, P));
        s.left_count = RoundIntD(static_cast<double>(lhi) * cnt_factor);
        s.left_sum_g = blg; s.left_sum_h = blh; s.left_int_g = lgi; s.left_int_h = lhi;
        s.right_output = (MC ? LeafOutputX(brg, brh, P, mc) : LeafOutputD(brg, brh, P));
        s.right_count = RoundIntD(static_cast<double>(rhi) * cnt_factor);
        s.right_sum_g = brg; s.right_sum_h = brh; s.right_int_g = rgi; s.right_int_h = rhi;
      } else {
        double best_lg, best_lh; int left_count;
        if (reverse) {
          const double rg = static_cast<double>(c.sg) * gscale, rh = KEPS + static_cast<double>(c.sh) * hscale;
          best_lg = sum_gradient - rg; best_lh = sum_hessian - rh; left_count = num_data - c.cnt;
        } else {
          best_lg = static_cast<double>(c.sg) * gscale; best_lh = KEPS + static_cast<double>(c.sh) * hscale; left_count = c.cnt;
        }
        s.threshold = reverse ? c.b - 1 : c.b;
        s.left_output = (MC ? LeafOutputX(best_lg, best_lh, P, mc) : LeafOutputD(best_lg, best_lh, P));
        s.left_count = left_count;
        s.left_sum_g = best_lg; s.left_sum_h = best_lh - KEPS;
        s.right_output = (MC ? LeafOutputX(sum_gradient - best_lg, sum_hessian - best_lh, P, mc) : LeafOutputD(sum_gradient - best_lg, sum_hessian - best_lh, P));
        s.right_count = num_data - left_count;
        s.right_sum_g = sum_gradient - best_lg; s.right_sum_h = sum_hessian - best_lh - KEPS;
      }
      s.gain = c.gain - min_gain_shift;
      s.default_left = reverse ? 1 : 0;
    }
    if (!two_scans && fm.missing_type == kMissingNaN) s.default_left = 0;
    if (MC && mc.mono != 0) { s.monotone_type = mc.mono; s.gain *= MonotonePenaltyD(depth, P.monotone_penalty); }
    *out = s;
    *splittable_out = any ? 1 : 0;
  }
}

// ------------------------------------------------------------------------------------------------ K6, categorical features
// One 256-thread CTA searches one categorical feature of one leaf, one bin per thread (FindBestThresholdCategoricalInner /
// ...IntInner).  Bin 0 (NaN, negative, rare and unseen categories) is never a candidate and always goes right.
//   few bins (num_bin <= max_cat_to_onehot): every category alone against the rest, one candidate per thread, block arg-max
//     (ties: the smaller bin, as the sequential loop's strict '>');
//   otherwise: the bins with at least cat_smooth rows are ranked by sum_gradient / (sum_hessian + cat_smooth) — every thread counts
//     the bins that precede its own (ties: the smaller bin first, the stable sort upstream) — and the two prefix scans over that
//     order, from the front and from the back (at most max_cat_threshold categories, groups of at least min_data_per_group rows,
//     cat_l2 added to lambda_l2), run sequentially on two threads of different warps: their `continue` / `break` rules carry state
//     (the size of the current group) from one position to the next, and there are at most 2 x max_cat_threshold positions.
// QUANT: every sum is int_sum * scale, without the kEpsilon terms of the float form.  Result valid in thread 0.  Uses one set of
// shared arrays: sub-block 0 only.
struct CatDirBest { double gain; int thr, left_count; long long ig, ih; double lg, lh; };

template <bool QUANT>
__device__ void BlockFindBestCat(long long g_in, long long h_in, const FeatureMetaDev fm, int f, bool root, long long tot_g,
                                 long long tot_h, double leaf_sum_g, double leaf_sum_h, int num_data, const LearnerParams& P,
                                 double gscale, double hscale, SplitDev* out, unsigned char* splittable_out,
                                 long long* out_tot_g, long long* out_tot_h, int sub, const McCtx mc = McCtx{0, 0, 0.0, 0.0}, int depth = 0) {
  __shared__ long long s_g[8], s_h[8];
  __shared__ double s_key[kFindThreads];
  __shared__ long long s_bg[kFindThreads], s_bh[kFindThreads];
  __shared__ unsigned char s_usedf[kFindThreads], s_sorted[kFindThreads];
  __shared__ Cand s_cand[8];
  __shared__ int s_wused[8];
  __shared__ unsigned int s_bits[8];
  __shared__ CatDirBest s_dir[2];
  __shared__ int s_pick_thr, s_pick_dir;
  const int b = threadIdx.x & (kFindThreads - 1), lane = b & 31, warp = b >> 5;
  const int num_bin = fm.num_bin;
  if (b < 8) s_bits[b] = 0u;
  if (root) {      // leaf totals = sum of all bins of this feature
    long long tg = g_in, th = h_in;
    for (int o = 16; o > 0; o >>= 1) { tg += __shfl_xor_sync(0xffffffffu, tg, o); th += __shfl_xor_sync(0xffffffffu, th, o); }
    if (lane == 0) { s_g[warp] = tg; s_h[warp] = th; }
    SubSync(sub);
    tot_g = 0; tot_h = 0;
    #pragma unroll
    for (int w = 0; w < 8; ++w) { tot_g += s_g[w]; tot_h += s_h[w]; }
    leaf_sum_g = static_cast<double>(tot_g) * gscale;
    leaf_sum_h = static_cast<double>(tot_h) * hscale;
    SubSync(sub);
  }
  *out_tot_g = tot_g; *out_tot_h = tot_h;
  double sum_gradient, sum_hessian, cnt_factor;
  if (QUANT) {
    sum_gradient = static_cast<double>(tot_g) * gscale;
    sum_hessian = static_cast<double>(tot_h) * hscale;
    cnt_factor = static_cast<double>(num_data) / static_cast<double>(tot_h);
  } else {
    sum_gradient = leaf_sum_g;
    sum_hessian = leaf_sum_h + 2 * KEPS;
    cnt_factor = num_data / sum_hessian;
  }
  const double eps = QUANT ? 0.0 : KEPS;
  const double gain_shift = LeafGainD(sum_gradient, sum_hessian, P);
  const double min_gain_shift = gain_shift + P.min_gain_to_split;
  const bool valid = b >= 1 && b < num_bin && !(QUANT && tot_h <= 0);
  const double grad = static_cast<double>(g_in) * gscale, hess = static_cast<double>(h_in) * hscale;
  const int cnt = !valid ? 0 : QUANT ? RoundIntD(static_cast<double>(h_in) * cnt_factor) : RoundIntD(hess * cnt_factor);
  const bool onehot = num_bin <= P.max_cat_to_onehot;
  SplitDev s = EmptySplit(f, fm.real_index);
  s.missing_type = fm.missing_type;
  s.default_left = 0;
  bool any = false;
  if (onehot) {
    Cand c; c.gain = KMINSCORE; c.b = -1; c.sg = 0; c.sh = 0; c.cnt = 0;
    if (valid) {
      bool ok = !(cnt < P.min_data_in_leaf || hess < P.min_sum_hessian) && !(num_data - cnt < P.min_data_in_leaf);
      double soh, sog;
      if (QUANT) { soh = static_cast<double>(tot_h - h_in) * hscale; sog = static_cast<double>(tot_g - g_in) * gscale; }
      else { soh = sum_hessian - hess - KEPS; sog = sum_gradient - grad; }
      ok = ok && !(soh < P.min_sum_hessian);
      if (ok) {
        const double gain = SplitGainX(sog, soh, grad, hess + eps, P, mc);
        if (gain > min_gain_shift) { c.gain = gain; c.b = b; c.sg = g_in; c.sh = h_in; c.cnt = cnt; }
      }
    }
    const Cand best = BlockBestCand(c, false, s_cand, sub);
    if (b == 0 && best.b >= 0) {
      any = true;
      const double blg = static_cast<double>(best.sg) * gscale, blh = static_cast<double>(best.sh) * hscale + eps;
      s.left_count = best.cnt; s.right_count = num_data - best.cnt;
      s.left_output = LeafOutputX(blg, blh, P, mc);
      if (QUANT) {
        const long long rgi = tot_g - best.sg, rhi = tot_h - best.sh;
        const double brg = static_cast<double>(rgi) * gscale, brh = static_cast<double>(rhi) * hscale;
        s.left_sum_g = blg; s.left_sum_h = blh; s.left_int_g = best.sg; s.left_int_h = best.sh;
        s.right_output = LeafOutputX(brg, brh, P, mc);
        s.right_sum_g = brg; s.right_sum_h = brh; s.right_int_g = rgi; s.right_int_h = rhi;
      } else {
        s.left_sum_g = blg; s.left_sum_h = blh - KEPS;
        s.right_output = LeafOutputX(sum_gradient - blg, sum_hessian - blh, P, mc);
        s.right_sum_g = sum_gradient - blg; s.right_sum_h = sum_hessian - blh - KEPS;
      }
      s.gain = best.gain - min_gain_shift;
      s.cat_bits[best.b >> 5] = 1u << (best.b & 31);
    }
  } else {
    LearnerParams Pc = P;
    Pc.l2 += P.cat_l2;
    const bool used = valid && static_cast<double>(cnt) >= P.cat_smooth;
    const double key = grad / (hess + P.cat_smooth);
    s_key[b] = key; s_bg[b] = g_in; s_bh[b] = h_in; s_usedf[b] = used ? 1 : 0;
    const unsigned um = __ballot_sync(0xffffffffu, used);
    if (lane == 0) s_wused[warp] = __popc(um);
    SubSync(sub);
    int used_bin = 0;
    #pragma unroll
    for (int w = 0; w < 8; ++w) used_bin += s_wused[w];
    if (used) {
      int r = 0;
      for (int j = 1; j < num_bin; ++j) {
        const double kj = s_key[j];
        if (s_usedf[j] && (kj < key || (kj == key && j < b))) ++r;
      }
      s_sorted[r] = static_cast<unsigned char>(b);
    }
    SubSync(sub);
    if (b == 0 || b == 32) {
      const int pass = b == 0 ? 0 : 1, dir = pass == 0 ? 1 : -1;
      int start_pos = pass == 0 ? 0 : used_bin - 1;
      const int max_num_cat = min(P.max_cat_threshold, (used_bin + 1) / 2);
      CatDirBest best; best.gain = KMINSCORE; best.thr = -1; best.left_count = 0; best.ig = best.ih = 0; best.lg = best.lh = 0.0;
      int cnt_cur_group = 0, left_count = 0;
      double sum_left_gradient = 0.0, sum_left_hessian = KEPS;
      long long ilg = 0, ilh = 0;
      for (int i = 0; i < used_bin && i < max_num_cat; ++i) {
        const int t = s_sorted[start_pos];
        start_pos += dir;
        const long long tg = s_bg[t], th = s_bh[t];
        ilg += tg; ilh += th;
        int c1;
        if (QUANT) {
          c1 = RoundIntD(static_cast<double>(th) * cnt_factor);
          sum_left_gradient = static_cast<double>(ilg) * gscale;
          sum_left_hessian = static_cast<double>(ilh) * hscale;
        } else {
          const double hb = static_cast<double>(th) * hscale;
          c1 = RoundIntD(hb * cnt_factor);
          sum_left_gradient += static_cast<double>(tg) * gscale;
          sum_left_hessian += hb;
        }
        left_count += c1;
        cnt_cur_group += c1;
        if (left_count < P.min_data_in_leaf || sum_left_hessian < P.min_sum_hessian) continue;
        const int right_count = num_data - left_count;
        if (right_count < P.min_data_in_leaf || right_count < P.min_data_per_group) break;
        const double sum_right_hessian = QUANT ? static_cast<double>(tot_h - ilh) * hscale : sum_hessian - sum_left_hessian;
        if (sum_right_hessian < P.min_sum_hessian) break;
        if (cnt_cur_group < P.min_data_per_group) continue;
        cnt_cur_group = 0;
        const double sum_right_gradient = QUANT ? static_cast<double>(tot_g - ilg) * gscale : sum_gradient - sum_left_gradient;
        const double gain = SplitGainX(sum_left_gradient, sum_left_hessian, sum_right_gradient, sum_right_hessian, Pc, mc);
        if (gain <= min_gain_shift) continue;
        if (gain > best.gain) {
          best.gain = gain; best.thr = i; best.left_count = left_count; best.ig = ilg; best.ih = ilh;
          best.lg = sum_left_gradient; best.lh = sum_left_hessian;
        }
      }
      s_dir[pass] = best;
    }
    SubSync(sub);
    if (b == 0) {
      int dir = 1;
      CatDirBest best = s_dir[0];
      if (s_dir[1].thr >= 0 && (best.thr < 0 || s_dir[1].gain > best.gain)) { best = s_dir[1]; dir = -1; }
      s_pick_thr = best.thr; s_pick_dir = dir;
      if (best.thr >= 0) {
        any = true;
        s.left_count = best.left_count; s.right_count = num_data - best.left_count;
        s.left_output = LeafOutputX(best.lg, best.lh, Pc, mc);
        if (QUANT) {
          const long long rgi = tot_g - best.ig, rhi = tot_h - best.ih;
          const double brg = static_cast<double>(rgi) * gscale, brh = static_cast<double>(rhi) * hscale;
          s.left_sum_g = best.lg; s.left_sum_h = best.lh; s.left_int_g = best.ig; s.left_int_h = best.ih;
          s.right_output = LeafOutputX(brg, brh, Pc, mc);
          s.right_sum_g = brg; s.right_sum_h = brh; s.right_int_g = rgi; s.right_int_h = rhi;
        } else {
          s.left_sum_g = best.lg; s.left_sum_h = best.lh - KEPS;
          s.right_output = LeafOutputX(sum_gradient - best.lg, sum_hessian - best.lh, Pc, mc);
          s.right_sum_g = sum_gradient - best.lg; s.right_sum_h = sum_hessian - best.lh - KEPS;
        }
        s.gain = best.gain - min_gain_shift;
      }
    }
    SubSync(sub);
    const int thr = s_pick_thr;
    if (b <= thr) {      // the categories of the chosen prefix
      const int t = s_pick_dir == 1 ? s_sorted[b] : s_sorted[used_bin - 1 - b];
      atomicOr(&s_bits[t >> 5], 1u << (t & 31));
    }
    SubSync(sub);
    if (b == 0 && any) {
      #pragma unroll
      for (int i = 0; i < 8; ++i) s.cat_bits[i] = s_bits[i];
    }
  }
  if (b == 0) {
    s.is_cat = 1;
    *out = s;
    *splittable_out = any ? 1 : 0;
  }
}

// ------------------------------------------------------------------------------------------------ K5 + K6, wide features
// A feature with more than 256 bins lives in S uint8 columns (dataset.h: column c codes the logical bins [253 c, 253 c + 253) as the
// bytes 1..253).  One 256-thread CTA — the unit of the feature's first column — searches it whole: the same scans, candidates and
// tie rules as BlockFindBest, over the logical bins, column after column with the prefix sums carried from one column to the next.
//   pass A  K5 for every column (child_bin(col, true): load / subtract from the parent / store) and the sums the scans need;
//   pass B  per column: block-wide inclusive scan + carry, one reverse and one forward candidate per thread, each thread keeping
//           the better of its candidates (reverse: the later column on equal gain, forward: the earlier).
// The result is handed on in column form: the split's column, the byte threshold in it and the byte whose rows follow
// default_left there, so that the partition and score kernels need not know the feature is wide.
template <bool QUANT, typename Load>
__device__ void BlockFindBestWide(Load child_bin, const FeatureMetaDev fm, int f, bool root, long long tot_g, long long tot_h,
                                  double leaf_sum_g, double leaf_sum_h, int num_data, const LearnerParams& P, double gscale,
                                  double hscale, SplitDev* out, unsigned char* splittable_out, long long* out_tot_g,
                                  long long* out_tot_h, const long long* slot, int sub, const McCtx mc = McCtx{0, 0, 0.0, 0.0}, int depth = 0) {
  __shared__ long long s_g[8], s_h[8];
  __shared__ int s_c[8];
  __shared__ Cand s_cand[8];
  __shared__ int s_flag;
  __shared__ long long s_bin0[2];
  __shared__ int s_bin0c;
  const int b = threadIdx.x & (kFindThreads - 1), lane = b & 31, warp = b >> 5;
  const int S = fm.wide_cols, NB = fm.num_bin;
  const bool zero_type = fm.missing_type == kMissingZero;
  const bool two_scans = fm.missing_type == kMissingNaN || zero_type;
  const int nan_bin = !two_scans ? -1 : zero_type ? fm.default_bin : NB - 1;      // the bin both scans leave out
  const int fstart = (two_scans && zero_type && fm.most_freq_bin == 0) ? 1 : 0;
  const bool in_col = b >= 1 && b <= kWideLocalBins;
  if (b == 0) s_flag = 0;
  if (b == 1) { s_bin0[0] = s_bin0[1] = 0; s_bin0c = 0; }      // thread 1 owns logical bin 0: it alone writes these
  auto block_sum3 = [&](long long& x, long long& y, int& z) {
    for (int o = 16; o > 0; o >>= 1) {
      x += __shfl_xor_sync(0xffffffffu, x, o); y += __shfl_xor_sync(0xffffffffu, y, o); z += __shfl_xor_sync(0xffffffffu, z, o);
    }
    SubSync(sub);
    if (lane == 0) { s_g[warp] = x; s_h[warp] = y; s_c[warp] = z; }
    SubSync(sub);
    x = 0; y = 0; z = 0;
    #pragma unroll
    for (int w = 0; w < 8; ++w) { x += s_g[w]; y += s_h[w]; z += s_c[w]; }
  };
  // ---- pass A: K5 for all columns; sums over all bins (the root's totals) and over the scanned bins
  long long ag = 0, ah = 0, ug = 0, uh = 0;
  for (int c = 0; c < S; ++c) {
    const longlong2 v = child_bin(f + c, true);
    const int L = c * kWideLocalBins + b - 1;
    if (in_col && L < NB) {
      ag += v.x; ah += v.y;
      if (L != nan_bin) {
        ug += v.x; uh += v.y;
        if (L == 0) { s_bin0[0] = v.x; s_bin0[1] = v.y; }
      }
    }
  }
  { int z = 0; block_sum3(ag, ah, z); }
  if (root) {
    tot_g = ag; tot_h = ah;
    leaf_sum_g = static_cast<double>(tot_g) * gscale;
    leaf_sum_h = static_cast<double>(tot_h) * hscale;
  }
  *out_tot_g = tot_g; *out_tot_h = tot_h;
  double sum_gradient, sum_hessian, cnt_factor;
  if (QUANT) {
    sum_gradient = static_cast<double>(tot_g) * gscale;
    sum_hessian = static_cast<double>(tot_h) * hscale;
    cnt_factor = static_cast<double>(num_data) / static_cast<double>(tot_h);
  } else {
    sum_gradient = leaf_sum_g;
    sum_hessian = leaf_sum_h + 2 * KEPS;
    cnt_factor = num_data / sum_hessian;
  }
  // float mode: the counts are sums of per-bin rounded counts, so their total over the scanned bins needs a pass of its own
  int uc = 0;
  if (!QUANT) {
    for (int c = 0; c < S; ++c) {
      const int L = c * kWideLocalBins + b - 1;
      if (in_col && L < NB && L != nan_bin) {
        const long long hv = slot[(static_cast<size_t>(f + c) * kBins + b) * 2 + 1];
        const int cv = RoundIntD(static_cast<double>(hv) * hscale * cnt_factor);
        uc += cv;
        if (L == 0) s_bin0c = cv;
      }
    }
  }
  block_sum3(ug, uh, uc);
  const long long Tg = ug, Th = uh; const int Tc = uc;
  const long long g0 = s_bin0[0], h0 = s_bin0[1]; const int c0 = s_bin0c;
  const double gain_shift = LeafGainD(sum_gradient, sum_hessian, P);
  const double min_gain_shift = gain_shift + P.min_gain_to_split;
  const int top = NB - 1 - ((two_scans && !zero_type) ? 1 : 0);
  bool splittable = false;
  Cand cr; cr.gain = KMINSCORE; cr.b = -1; cr.sg = 0; cr.sh = 0; cr.cnt = 0;
  Cand cf = cr;
  long long Cg = 0, Ch = 0; int Cc = 0;      // sums of the scanned bins of the columns before this one
  // ---- pass B
  for (int c = 0; c < S; ++c) {
    const int L = c * kWideLocalBins + b - 1;
    const bool use = in_col && L < NB && L != nan_bin;
    long long gv = 0, hv = 0; int cv = 0;
    if (use) {
      const longlong2 t = *reinterpret_cast<const longlong2*>(slot + (static_cast<size_t>(f + c) * kBins + b) * 2);
      gv = t.x; hv = t.y;
      if (!QUANT) cv = RoundIntD(static_cast<double>(hv) * hscale * cnt_factor);
    }
    long long pg = gv, ph = hv; int pc = cv;
    #pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const long long tg = __shfl_up_sync(0xffffffffu, pg, o);
      const long long th = __shfl_up_sync(0xffffffffu, ph, o);
      const int tc = __shfl_up_sync(0xffffffffu, pc, o);
      if (lane >= o) { pg += tg; ph += th; pc += tc; }
    }
    SubSync(sub);
    if (lane == 31) { s_g[warp] = pg; s_h[warp] = ph; s_c[warp] = pc; }
    SubSync(sub);
    long long colg = 0, colh = 0; int colc = 0;
    #pragma unroll
    for (int w = 0; w < 8; ++w) {
      if (w < warp) { pg += s_g[w]; ph += s_h[w]; pc += s_c[w]; }
      colg += s_g[w]; colh += s_h[w]; colc += s_c[w];
    }
    pg += Cg; ph += Ch; pc += Cc;
    // reverse scan candidate: right = bins [L, top], threshold = L - 1
    if (use && L >= 1 && L <= top) {
      const long long rgi = Tg - (pg - gv), rhi = Th - (ph - hv);
      double gain = 0.0; int rc; bool ok = true;
      if (QUANT) {
        rc = RoundIntD(static_cast<double>(rhi) * cnt_factor);
        const double srh = static_cast<double>(rhi) * hscale;
        const int lc = num_data - rc;
        const long long lgi = tot_g - rgi, lhi = tot_h - rhi;
        const double slh = static_cast<double>(lhi) * hscale;
        ok = !(rc < P.min_data_in_leaf || srh < P.min_sum_hessian) && !(lc < P.min_data_in_leaf) && !(slh < P.min_sum_hessian);
        if (ok) gain = SplitGainX(static_cast<double>(lgi) * gscale, slh, static_cast<double>(rgi) * gscale, srh, P, mc);
      } else {
        rc = Tc - (pc - cv);
        const double rh = KEPS + static_cast<double>(rhi) * hscale;
        const int lc = num_data - rc;
        const double lh = sum_hessian - rh;
        ok = !(rc < P.min_data_in_leaf || rh < P.min_sum_hessian) && !(lc < P.min_data_in_leaf) && !(lh < P.min_sum_hessian);
        if (ok) { const double rg = static_cast<double>(rgi) * gscale; gain = SplitGainX(sum_gradient - rg, lh, rg, rh, P, mc); }
      }
      if (ok && gain > min_gain_shift) {
        splittable = true;
        if (gain >= cr.gain) { cr.gain = gain; cr.b = L; cr.sg = rgi; cr.sh = rhi; cr.cnt = rc; }      // a later column: the larger bin
      }
    }
    // forward scan candidate: left = bins [fstart, L], threshold = L
    if (two_scans && use && L >= fstart && L <= NB - 2) {
      const long long lgi = pg - (fstart ? g0 : 0), lhi = ph - (fstart ? h0 : 0);
      double gain = 0.0; int lc; bool ok = true;
      if (QUANT) {
        lc = RoundIntD(static_cast<double>(lhi) * cnt_factor);
        const double slh = static_cast<double>(lhi) * hscale;
        const int rc = num_data - lc;
        const long long rgi = tot_g - lgi, rhi = tot_h - lhi;
        const double srh = static_cast<double>(rhi) * hscale;
        ok = !(lc < P.min_data_in_leaf || slh < P.min_sum_hessian) && !(rc < P.min_data_in_leaf) && !(srh < P.min_sum_hessian);
        if (ok) gain = SplitGainX(static_cast<double>(lgi) * gscale, slh, static_cast<double>(rgi) * gscale, srh, P, mc);
      } else {
        lc = pc - (fstart ? c0 : 0);
        const double lh = KEPS + static_cast<double>(lhi) * hscale;
        const int rc = num_data - lc;
        const double rh = sum_hessian - lh;
        ok = !(lc < P.min_data_in_leaf || lh < P.min_sum_hessian) && !(rc < P.min_data_in_leaf) && !(rh < P.min_sum_hessian);
        if (ok) { const double lg = static_cast<double>(lgi) * gscale; gain = SplitGainX(lg, lh, sum_gradient - lg, rh, P, mc); }
      }
      if (ok && gain > min_gain_shift) {
        splittable = true;
        if (gain > cf.gain) { cf.gain = gain; cf.b = L; cf.sg = lgi; cf.sh = lhi; cf.cnt = lc; }        // an earlier column keeps an equal gain
      }
    }
    Cg += colg; Ch += colh; Cc += colc;
  }
  if (splittable) s_flag = 1;
  const Cand best_r = BlockBestCand(cr, true, s_cand, sub);
  Cand best_f = cf;
  if (two_scans) best_f = BlockBestCand(cf, false, s_cand, sub);
  SubSync(sub);
  if (b == 0) {
    const bool any = s_flag != 0;
    const bool accept = !(mc.on && mc.cmin > mc.cmax);      // an empty output range accepts no candidate (the feature stays "splittable")
    SplitDev s = EmptySplit(f, fm.real_index);
    s.missing_type = fm.missing_type;
    int thr = 0;
    for (int pass = 0; pass < 2; ++pass) {
      const Cand& c = pass == 0 ? best_r : best_f;
      if (pass == 1 && !two_scans) break;
      if (!(any && accept && c.b >= 0 && c.gain > s.gain + min_gain_shift)) continue;
      const bool reverse = pass == 0;
      if (QUANT) {
        const long long lgi = reverse ? tot_g - c.sg : c.sg, lhi = reverse ? tot_h - c.sh : c.sh;
        const long long rgi = tot_g - lgi, rhi = tot_h - lhi;
        const double blg = static_cast<double>(lgi) * gscale, blh = static_cast<double>(lhi) * hscale;
        const double brg = static_cast<double>(rgi) * gscale, brh = static_cast<double>(rhi) * hscale;
        s.left_output = LeafOutputX(blg, blh, P, mc);
        s.left_count = RoundIntD(static_cast<double>(lhi) * cnt_factor);
        s.left_sum_g = blg; s.left_sum_h = blh; s.left_int_g = lgi; s.left_int_h = lhi;
        s.right_output = LeafOutputX(brg, brh, P, mc);
        s.right_count = RoundIntD(static_cast<double>(rhi) * cnt_factor);
        s.right_sum_g = brg; s.right_sum_h = brh; s.right_int_g = rgi; s.right_int_h = rhi;
      } else {
        double best_lg, best_lh; int left_count;
        if (reverse) {
          const double rg = static_cast<double>(c.sg) * gscale, rh = KEPS + static_cast<double>(c.sh) * hscale;
          best_lg = sum_gradient - rg; best_lh = sum_hessian - rh; left_count = num_data - c.cnt;
        } else {
          best_lg = static_cast<double>(c.sg) * gscale; best_lh = KEPS + static_cast<double>(c.sh) * hscale; left_count = c.cnt;
        }
        s.left_output = LeafOutputX(best_lg, best_lh, P, mc);
        s.left_count = left_count;
        s.left_sum_g = best_lg; s.left_sum_h = best_lh - KEPS;
        s.right_output = LeafOutputX(sum_gradient - best_lg, sum_hessian - best_lh, P, mc);
        s.right_count = num_data - left_count;
        s.right_sum_g = sum_gradient - best_lg; s.right_sum_h = sum_hessian - best_lh - KEPS;
      }
      thr = reverse ? c.b - 1 : c.b;
      s.gain = c.gain - min_gain_shift;
      s.default_left = reverse ? 1 : 0;
    }
    if (!two_scans && fm.missing_type == kMissingNaN) s.default_left = 0;
    if (mc.on && mc.mono != 0) { s.monotone_type = mc.mono; s.gain *= MonotonePenaltyD(depth, P.monotone_penalty); }
    // column form: the split's column, the byte threshold in it, and the byte whose rows follow default_left there
    const int ct = thr / kWideLocalBins;
    const int miss = fm.missing_type == kMissingNaN ? NB - 1 : zero_type ? fm.default_bin : -1;
    s.feature = f + ct;
    s.threshold = thr % kWideLocalBins + 1;
    s.nan_bin = miss < 0 ? -1 : (miss / kWideLocalBins == ct ? miss % kWideLocalBins + 1 : kWideMissing);
    *out = s;
    *splittable_out = any ? 1 : 0;
  }
}

// Leaf totals of the root seen from a further column of a wide feature (every unit of the root reports them): the sum over the
// feature's bins, read without storing anything.
template <typename Load>
__device__ void WideRootTotals(Load child_bin, const FeatureMetaDev first, int f0, long long* out_tot_g, long long* out_tot_h, int sub) {
  __shared__ long long s_g[8], s_h[8];
  const int b = threadIdx.x & (kFindThreads - 1), lane = b & 31, warp = b >> 5;
  long long ag = 0, ah = 0;
  for (int c = 0; c < first.wide_cols; ++c) {
    const longlong2 v = child_bin(f0 + c, false);
    const int L = c * kWideLocalBins + b - 1;
    if (b >= 1 && b <= kWideLocalBins && L < first.num_bin) { ag += v.x; ah += v.y; }
  }
  for (int o = 16; o > 0; o >>= 1) { ag += __shfl_xor_sync(0xffffffffu, ag, o); ah += __shfl_xor_sync(0xffffffffu, ah, o); }
  SubSync(sub);
  if (lane == 0) { s_g[warp] = ag; s_h[warp] = ah; }
  SubSync(sub);
  ag = 0; ah = 0;
  #pragma unroll
  for (int w = 0; w < 8; ++w) { ag += s_g[w]; ah += s_h[w]; }
  *out_tot_g = ag; *out_tot_h = ah;
}

// ------------------------------------------------------------------------------------------------ K6 as a warp scan
// One WARP searches one feature of one leaf: lane l owns bins [8 l, 8 l + 8) — eight 16-byte histogram entries, one coalesced
// 4 KB read per warp — sums them, the lane totals are prefix-summed with shuffles, and every lane evaluates its eight thresholds
// (the same expressions, in the same order, as the sequential scan: FindBestThresholdSequentially / ...Int with a reverse
// scan, plus a forward scan when the feature has a NaN bin and more than two bins).  The arg-max over thresholds keeps the
// sequential scan's tie rules (reverse: the larger bin on equal gain, forward: the smaller).  No shared memory, no block
// barrier: 2000 features x 2 children x 8 leaves are 32000 warps (3-4 waves) instead of 32000 256-thread CTAs (27 waves).
// g_in / h_in: the lane's eight bins of the leaf's histogram.  The result is written by lane 0.
constexpr int kBinsPerLane = kBins / 32;

__device__ __forceinline__ bool CandBetter(const Cand& t, const Cand& cur, bool prefer_larger_b) {
  return (t.gain > cur.gain) || (t.gain == cur.gain && t.b >= 0 && (cur.b < 0 || (prefer_larger_b ? t.b > cur.b : t.b < cur.b)));
}
__device__ __forceinline__ Cand WarpBestCand(Cand c, bool prefer_larger_b) {
  for (int o = 16; o > 0; o >>= 1) {
    Cand t;
    t.gain = __shfl_xor_sync(0xffffffffu, c.gain, o);
    t.b = __shfl_xor_sync(0xffffffffu, c.b, o);
    t.sg = __shfl_xor_sync(0xffffffffu, c.sg, o);
    t.sh = __shfl_xor_sync(0xffffffffu, c.sh, o);
    t.cnt = __shfl_xor_sync(0xffffffffu, c.cnt, o);
    if (CandBetter(t, c, prefer_larger_b)) c = t;
  }
  return c;
}

template <bool QUANT>
__device__ __forceinline__ void WarpFindBest(const long long (&g_in)[kBinsPerLane], const long long (&h_in)[kBinsPerLane], const FeatureMetaDev fm,
                                             int f, bool root, long long tot_g, long long tot_h, double leaf_sum_g, double leaf_sum_h,
                                             int num_data, const LearnerParams& P, double gscale, double hscale, SplitDev* out,
                                             unsigned char* splittable_out, long long* out_tot_g, long long* out_tot_h) {
  const int lane = threadIdx.x & 31;
  const int num_bin = fm.num_bin;
  const bool zero_type = fm.missing_type == kMissingZero;      // (see BlockFindBest)
  const bool two_scans = (num_bin > 2 && (fm.missing_type == kMissingNaN || zero_type));
  const int nan_bin = !two_scans ? -1 : zero_type ? fm.default_bin : num_bin - 1;
  const int fstart = (two_scans && zero_type && fm.most_freq_bin == 0) ? 1 : 0;
  if (root) {      // the leaf totals = the sum of all bins of this feature
    long long tg = 0, th = 0;
    #pragma unroll
    for (int j = 0; j < kBinsPerLane; ++j) { tg += g_in[j]; th += h_in[j]; }
    for (int o = 16; o > 0; o >>= 1) { tg += __shfl_xor_sync(0xffffffffu, tg, o); th += __shfl_xor_sync(0xffffffffu, th, o); }
    tot_g = tg; tot_h = th;
    leaf_sum_g = static_cast<double>(tot_g) * gscale;
    leaf_sum_h = static_cast<double>(tot_h) * hscale;
  }
  *out_tot_g = tot_g; *out_tot_h = tot_h;
  double sum_gradient, sum_hessian, cnt_factor;
  if (QUANT) {
    sum_gradient = static_cast<double>(tot_g) * gscale;
    sum_hessian = static_cast<double>(tot_h) * hscale;
    cnt_factor = static_cast<double>(num_data) / static_cast<double>(tot_h);
  } else {
    sum_gradient = leaf_sum_g;
    sum_hessian = leaf_sum_h + 2 * KEPS;
    cnt_factor = num_data / sum_hessian;
  }
  // the lane's totals over its (masked) bins; the bins are walked again below with a running prefix, so only the raw bins stay live
  auto masked = [&](int j, long long* g, long long* h, int* cnt) {
    const int b = lane * kBinsPerLane + j;
    const bool use = b < num_bin && b != nan_bin;
    *g = use ? g_in[j] : 0; *h = use ? h_in[j] : 0;
    *cnt = (!QUANT && use) ? RoundIntD(static_cast<double>(*h) * hscale * cnt_factor) : 0;
  };
  long long lg = 0, lh = 0; int lc = 0;
  #pragma unroll
  for (int j = 0; j < kBinsPerLane; ++j) {
    long long g, h; int cn;
    masked(j, &g, &h, &cn);
    lg += g; lh += h; lc += cn;
  }
  // exclusive prefix of the lane totals over the warp, and the totals
  long long og = lg, oh = lh; int oc = lc;
  #pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const long long tg = __shfl_up_sync(0xffffffffu, og, o);
    const long long th = __shfl_up_sync(0xffffffffu, oh, o);
    const int tc = __shfl_up_sync(0xffffffffu, oc, o);
    if (lane >= o) { og += tg; oh += th; oc += tc; }
  }
  const long long Tg = __shfl_sync(0xffffffffu, og, 31), Th = __shfl_sync(0xffffffffu, oh, 31);
  const int Tc = __shfl_sync(0xffffffffu, oc, 31);
  og -= lg; oh -= lh; oc -= lc;      // what the lower lanes hold
  // bin 0 (lane 0's first bin), which a forward scan that starts at bin 1 leaves out
  long long g0, h0; int c0;
  masked(0, &g0, &h0, &c0);
  g0 = __shfl_sync(0xffffffffu, g0, 0); h0 = __shfl_sync(0xffffffffu, h0, 0); c0 = __shfl_sync(0xffffffffu, c0, 0);
  if (!fstart) { g0 = 0; h0 = 0; c0 = 0; }
  const double gain_shift = LeafGainD(sum_gradient, sum_hessian, P);
  const double min_gain_shift = gain_shift + P.min_gain_to_split;
  bool splittable = false;
  const int top = num_bin - 1 - ((two_scans && !zero_type) ? 1 : 0);
  Cand cr; cr.gain = KMINSCORE; cr.b = -1; cr.sg = 0; cr.sh = 0; cr.cnt = 0;
  Cand cf = cr;
  #pragma unroll
  for (int j = 0; j < kBinsPerLane; ++j) {
    const int b = lane * kBinsPerLane + j;
    long long gvj, hvj; int cvj;
    masked(j, &gvj, &hvj, &cvj);
    og += gvj; oh += hvj; oc += cvj;
    const long long ipg = og, iph = oh;      // inclusive prefix over bins [0, b]
    const int ipc = oc;
    // ---- reverse scan candidate: right = bins [b, top], threshold = b - 1
    if (b >= 1 && b <= top && b != nan_bin) {
      const long long rgi = Tg - (ipg - gvj), rhi = Th - (iph - hvj);
      double gain = 0.0; int rc; bool ok = true;
      if (QUANT) {
        rc = RoundIntD(static_cast<double>(rhi) * cnt_factor);
        const double srh = static_cast<double>(rhi) * hscale;
        const int lcn = num_data - rc;
        const long long lgi = tot_g - rgi, lhi = tot_h - rhi;
        const double slh = static_cast<double>(lhi) * hscale;
        ok = !(rc < P.min_data_in_leaf || srh < P.min_sum_hessian) && !(lcn < P.min_data_in_leaf) && !(slh < P.min_sum_hessian);
        if (ok) gain = SplitGainD(static_cast<double>(lgi) * gscale, slh, static_cast<double>(rgi) * gscale, srh, P);
      } else {
        rc = Tc - (ipc - cvj);
        const double rh = KEPS + static_cast<double>(rhi) * hscale;
        const int lcn = num_data - rc;
        const double lhh = sum_hessian - rh;
        ok = !(rc < P.min_data_in_leaf || rh < P.min_sum_hessian) && !(lcn < P.min_data_in_leaf) && !(lhh < P.min_sum_hessian);
        if (ok) { const double rg = static_cast<double>(rgi) * gscale; gain = SplitGainD(sum_gradient - rg, lhh, rg, rh, P); }
      }
      if (ok && gain > min_gain_shift) {
        splittable = true;
        Cand t; t.gain = gain; t.b = b; t.sg = rgi; t.sh = rhi; t.cnt = rc;
        if (CandBetter(t, cr, true)) cr = t;
      }
    }
    // ---- forward scan candidate (NaN as missing): left = bins [0, b], threshold = b
    if (two_scans && b >= fstart && b <= num_bin - 2 && b != nan_bin) {
      const long long lgi = ipg - g0, lhi = iph - h0;
      double gain = 0.0; int lcn; bool ok = true;
      if (QUANT) {
        lcn = RoundIntD(static_cast<double>(lhi) * cnt_factor);
        const double slh = static_cast<double>(lhi) * hscale;
        const int rc = num_data - lcn;
        const long long rgi = tot_g - lgi, rhi = tot_h - lhi;
        const double srh = static_cast<double>(rhi) * hscale;
        ok = !(lcn < P.min_data_in_leaf || slh < P.min_sum_hessian) && !(rc < P.min_data_in_leaf) && !(srh < P.min_sum_hessian);
        if (ok) gain = SplitGainD(static_cast<double>(lgi) * gscale, slh, static_cast<double>(rgi) * gscale, srh, P);
      } else {
        lcn = ipc - c0;
        const double lhh = KEPS + static_cast<double>(lhi) * hscale;
        const int rc = num_data - lcn;
        const double rh = sum_hessian - lhh;
        ok = !(lcn < P.min_data_in_leaf || lhh < P.min_sum_hessian) && !(rc < P.min_data_in_leaf) && !(rh < P.min_sum_hessian);
        if (ok) { const double lgd = static_cast<double>(lgi) * gscale; gain = SplitGainD(lgd, lhh, sum_gradient - lgd, rh, P); }
      }
      if (ok && gain > min_gain_shift) {
        splittable = true;
        Cand t; t.gain = gain; t.b = b; t.sg = lgi; t.sh = lhi; t.cnt = lcn;
        if (CandBetter(t, cf, false)) cf = t;
      }
    }
  }
  const bool any = __any_sync(0xffffffffu, splittable);
  const Cand best_r = WarpBestCand(cr, true);
  Cand best_f = cf;
  if (two_scans) best_f = WarpBestCand(cf, false);
  if (lane == 0) {
    SplitDev s = EmptySplit(f, fm.real_index);
    s.missing_type = fm.missing_type;
    s.nan_bin = (fm.missing_type == kMissingNaN) ? fm.num_bin - 1 : zero_type ? fm.default_bin : -1;      // rows of this bin follow default_left
    for (int pass = 0; pass < 2; ++pass) {
      const Cand& c = pass == 0 ? best_r : best_f;
      if (pass == 1 && !two_scans) break;
      if (!(any && c.b >= 0 && c.gain > s.gain + min_gain_shift)) continue;
      const bool reverse = pass == 0;
      if (QUANT) {
        const long long lgi = reverse ? tot_g - c.sg : c.sg, lhi = reverse ? tot_h - c.sh : c.sh;
        const long long rgi = tot_g - lgi, rhi = tot_h - lhi;
        const double blg = static_cast<double>(lgi) * gscale, blh = static_cast<double>(lhi) * hscale;
        const double brg = static_cast<double>(rgi) * gscale, brh = static_cast<double>(rhi) * hscale;
        s.threshold = reverse ? c.b - 1 : c.b;
        s.left_output = LeafOutputD(blg, blh, P);
        s.left_count = RoundIntD(static_cast<double>(lhi) * cnt_factor);
        s.left_sum_g = blg; s.left_sum_h = blh; s.left_int_g = lgi; s.left_int_h = lhi;
        s.right_output = LeafOutputD(brg, brh, P);
        s.right_count = RoundIntD(static_cast<double>(rhi) * cnt_factor);
        s.right_sum_g = brg; s.right_sum_h = brh; s.right_int_g = rgi; s.right_int_h = rhi;
      } else {
        double best_lg, best_lh; int left_count;
        if (reverse) {
          const double rg = static_cast<double>(c.sg) * gscale, rh = KEPS + static_cast<double>(c.sh) * hscale;
          best_lg = sum_gradient - rg; best_lh = sum_hessian - rh; left_count = num_data - c.cnt;
        } else {
          best_lg = static_cast<double>(c.sg) * gscale; best_lh = KEPS + static_cast<double>(c.sh) * hscale; left_count = c.cnt;
        }
        s.threshold = reverse ? c.b - 1 : c.b;
        s.left_output = LeafOutputD(best_lg, best_lh, P);
        s.left_count = left_count;
        s.left_sum_g = best_lg; s.left_sum_h = best_lh - KEPS;
        s.right_output = LeafOutputD(sum_gradient - best_lg, sum_hessian - best_lh, P);
        s.right_count = num_data - left_count;
        s.right_sum_g = sum_gradient - best_lg; s.right_sum_h = sum_hessian - best_lh - KEPS;
      }
      s.gain = c.gain - min_gain_shift;
      s.default_left = reverse ? 1 : 0;
    }
    if (!two_scans && fm.missing_type == kMissingNaN) s.default_left = 0;
    *out = s;
    *splittable_out = any ? 1 : 0;
  }
}

// K7 (leaf partition) lives in batch_tree.cuh.

}  // namespace lgb
