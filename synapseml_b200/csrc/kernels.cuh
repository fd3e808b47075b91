// Device kernels of the training hot path (K3-K8).  Algorithm specification: SURVEY.md Appendix A.5-A.9 (upstream
// objective / feature_histogram / data_partition / gbdt); the kernels are designed for sm_100a from that specification.
//
// Conventions
//  * Histograms are exact integers in every mode.  Quantized mode accumulates the int8 (grad, hess) pairs; float mode
//    first converts float32 grad/hess to 19-bit fixed point with power-of-two scales (one scale per iteration from
//    max|g|, max|h|), so sums are order-independent: the result does not depend on the number of GPUs or on atomic
//    ordering, and the int64 global histogram converts back to float exactly (scale is 2^k).
//  * Global histogram entry = (int64 sum_g, int64 sum_h) per (feature, bin), 256 bin slots per feature.
//  * All per-tree control flow lives in device memory (TreeCtrl / LeafDev) so a whole tree is enqueued with no host
//    round trip; kernels return immediately once ctrl->done is set.
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
__global__ void k_grad_binary(const double* __restrict__ score, const float* __restrict__ label, const float* __restrict__ weight,
                              float* __restrict__ grad, float* __restrict__ hess, int n, double sigmoid, double lw_neg, double lw_pos) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const int is_pos = label[i] > 0.0f;
    const int lab = is_pos ? 1 : -1;
    const double lw = is_pos ? lw_pos : lw_neg;
    const double response = -lab * sigmoid / (1.0f + SoftExp(lab * sigmoid * score[i]));
    const double abs_response = fabs(response);
    if (weight == nullptr) {
      grad[i] = static_cast<float>(response * lw);
      hess[i] = static_cast<float>(abs_response * (sigmoid - abs_response) * lw);
    } else {
      grad[i] = static_cast<float>(response * lw * weight[i]);
      hess[i] = static_cast<float>(abs_response * (sigmoid - abs_response) * lw * weight[i]);
    }
  }
}

__global__ void k_grad_l2(const double* __restrict__ score, const float* __restrict__ label, const float* __restrict__ weight,
                          float* __restrict__ grad, float* __restrict__ hess, int n) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    if (weight == nullptr) {
      grad[i] = static_cast<float>(score[i] - label[i]);
      hess[i] = 1.0f;
    } else {
      grad[i] = static_cast<float>(static_cast<float>(score[i] - label[i]) * weight[i]);
      hess[i] = static_cast<float>(weight[i]);
    }
  }
}

// LambdaRank-NDCG.  One CTA per query.  Step 1: stable rank of every document by score (descending) by counting, which
// equals std::stable_sort's order.  Step 2: one thread per document replays exactly the pairs that touch it, in the
// CPU's (i, j) order, so its float32 lambda/hessian sums are bit-identical to the sequential loop.  Step 3: thread 0
// re-walks the pairs in (i, j) order for the double-precision sum_lambdas (needed for the log2 normalisation).
struct LambdarankArgs {
  const double* score; const float* label; const float* weight; const int* qb;
  const double* inv_max_dcg; const double* label_gain; const double* discount; const double* sigmoid_table;
  float* grad; float* hess;
  double sigmoid, min_sigmoid_input, max_sigmoid_input, idx_factor;
  int truncation_level, norm, table_size, max_cnt;
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

__global__ void k_grad_lambdarank(LambdarankArgs a) {
  extern __shared__ __align__(16) unsigned char lr_smem[];
  const int q = blockIdx.x;
  const int start = a.qb[q];
  const int cnt = a.qb[q + 1] - start;
  double* s_score = reinterpret_cast<double*>(lr_smem);                 // [max_cnt] sorted by rank
  float* s_label = reinterpret_cast<float*>(s_score + a.max_cnt);      // [max_cnt]
  int* s_doc = reinterpret_cast<int*>(s_label + a.max_cnt);            // rank -> doc
  __shared__ double s_norm;
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
  }
  __syncthreads();
  const double inverse_max_dcg = a.inv_max_dcg[q];
  const double best_score = s_score[0];
  int worst_idx = cnt - 1;
  if (worst_idx > 0 && s_score[worst_idx] == KMINSCORE) worst_idx -= 1;
  const double worst_score = s_score[worst_idx];
  const bool norm_on = a.norm && best_score != worst_score;
  const int T = a.truncation_level;
  const int imax = min(cnt - 1, T);      // i ranges over [0, imax)
  for (int r = threadIdx.x; r < cnt; r += blockDim.x) {
    float lam = 0.0f, hes = 0.0f;
    if (s_score[r] != KMINSCORE) {
      // as the later element j = r of pairs (i, r), i < min(r, imax)
      for (int i = 0; i < min(r, imax); ++i) {
        if (s_score[i] == KMINSCORE) continue;
        double pl, ph; bool i_high;
        if (!LrPair(a, s_score, s_label, i, r, inverse_max_dcg, norm_on, &pl, &ph, &i_high)) continue;
        if (i_high) lam -= static_cast<float>(pl); else lam += static_cast<float>(pl);   // r is low when i is high
        hes += static_cast<float>(ph);
      }
      // as the earlier element i = r of pairs (r, j), j > r
      if (r < imax) {
        for (int j = r + 1; j < cnt; ++j) {
          if (s_score[j] == KMINSCORE) continue;
          double pl, ph; bool i_high;
          if (!LrPair(a, s_score, s_label, r, j, inverse_max_dcg, norm_on, &pl, &ph, &i_high)) continue;
          if (i_high) lam += static_cast<float>(pl); else lam -= static_cast<float>(pl);
          hes += static_cast<float>(ph);
        }
      }
    }
    a.grad[start + s_doc[r]] = lam;
    a.hess[start + s_doc[r]] = hes;
  }
  if (a.norm) {
    if (threadIdx.x == 0) {
      double sum_lambdas = 0.0;
      for (int i = 0; i < imax; ++i) {
        if (s_score[i] == KMINSCORE) continue;
        for (int j = i + 1; j < cnt; ++j) {
          if (s_score[j] == KMINSCORE) continue;
          double pl, ph; bool i_high;
          if (!LrPair(a, s_score, s_label, i, j, inverse_max_dcg, norm_on, &pl, &ph, &i_high)) continue;
          sum_lambdas -= 2 * pl;
        }
      }
      s_norm = (sum_lambdas > 0) ? SoftLog2(1 + sum_lambdas) / sum_lambdas : -1.0;
    }
    __syncthreads();
    if (s_norm >= 0.0) {
      const double nf = s_norm;
      for (int d = threadIdx.x; d < cnt; d += blockDim.x) {
        a.grad[start + d] = static_cast<float>(a.grad[start + d] * nf);
        a.hess[start + d] = static_cast<float>(a.hess[start + d] * nf);
      }
    }
  }
  if (a.weight != nullptr) {
    __syncthreads();
    for (int d = threadIdx.x; d < cnt; d += blockDim.x) {
      a.grad[start + d] = static_cast<float>(a.grad[start + d] * a.weight[start + d]);
      a.hess[start + d] = static_cast<float>(a.hess[start + d] * a.weight[start + d]);
    }
  }
}

__global__ void k_add_constant(double* score, int n, double v) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) score[i] += v;
}

// max|g|, max|h| over rows (non-negative float bit patterns order like unsigned ints)
__global__ void k_absmax(const float* __restrict__ grad, const float* __restrict__ hess, int n, TreeCtrl* ctrl) {
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

__global__ void k_scales(TreeCtrl* ctrl, LearnerParams lp) {
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

__global__ void k_quantize(const float* __restrict__ grad, const float* __restrict__ hess, int n, const TreeCtrl* __restrict__ ctrl,
                           LearnerParams lp, unsigned int iter, unsigned int row_offset, uint16_t* __restrict__ gh16,
                           int2* __restrict__ ghfix) {
  const double inv_g = ctrl->inv_grad_scale, inv_h = ctrl->inv_hess_scale;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
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
  return s;
}

// Split search runs on 256-thread sub-blocks (one per feature x child); a CTA may hold more than one of them (the persistent
// tree kernel runs two), so they synchronise on named barriers instead of __syncthreads().
constexpr int kFindThreads = 256;
constexpr int kMaxFindSubs = 2;
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
template <bool QUANT>
__device__ void BlockFindBest(long long g_in, long long h_in, const FeatureMetaDev fm, int f, bool root, long long tot_g,
                              long long tot_h, double leaf_sum_g, double leaf_sum_h, int num_data, const LearnerParams& P,
                              double gscale, double hscale, SplitDev* out, unsigned char* splittable_out,
                              long long* out_tot_g, long long* out_tot_h, int sub) {
  __shared__ long long s_g_all[kMaxFindSubs][8], s_h_all[kMaxFindSubs][8];
  __shared__ int s_c_all[kMaxFindSubs][8];
  __shared__ Cand s_cand_all[kMaxFindSubs][8];
  __shared__ int s_flag_all[kMaxFindSubs];
  long long* s_g = s_g_all[sub]; long long* s_h = s_h_all[sub]; int* s_c = s_c_all[sub]; Cand* s_cand = s_cand_all[sub];
  int& s_flag = s_flag_all[sub];
  const int b = threadIdx.x & (kFindThreads - 1), lane = b & 31, warp = b >> 5;
  const int num_bin = fm.num_bin;
  const bool two_scans = (num_bin > 2 && fm.missing_type == kMissingNaN);
  const int nan_bin = two_scans ? num_bin - 1 : -1;
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
  const int top = num_bin - 1 - (two_scans ? 1 : 0);
  Cand cr; cr.gain = KMINSCORE; cr.b = -1; cr.sg = 0; cr.sh = 0; cr.cnt = 0;
  if (b >= 1 && b <= top) {
    const long long rgi = Tg - (pg - gv), rhi = Th - (ph - hv);
    double gain = 0.0; int rc; bool ok = true;
    if (QUANT) {
      rc = RoundIntD(static_cast<double>(rhi) * cnt_factor);
      const double srh = static_cast<double>(rhi) * hscale;
      const int lc = num_data - rc;
      const long long lgi = tot_g - rgi, lhi = tot_h - rhi;
      const double slh = static_cast<double>(lhi) * hscale;
      ok = !(rc < P.min_data_in_leaf || srh < P.min_sum_hessian) && !(lc < P.min_data_in_leaf) && !(slh < P.min_sum_hessian);
      if (ok) gain = SplitGainD(static_cast<double>(lgi) * gscale, slh, static_cast<double>(rgi) * gscale, srh, P);
    } else {
      rc = Tc - (pc - cv);
      const double rh = KEPS + static_cast<double>(rhi) * hscale;
      const int lc = num_data - rc;
      const double lh = sum_hessian - rh;
      ok = !(rc < P.min_data_in_leaf || rh < P.min_sum_hessian) && !(lc < P.min_data_in_leaf) && !(lh < P.min_sum_hessian);
      if (ok) { const double rg = static_cast<double>(rgi) * gscale; gain = SplitGainD(sum_gradient - rg, lh, rg, rh, P); }
    }
    if (ok && gain > min_gain_shift) { splittable = true; cr.gain = gain; cr.b = b; cr.sg = rgi; cr.sh = rhi; cr.cnt = rc; }
  }
  // ---- forward scan candidate (NaN as missing): left = bins [0, b], threshold = b
  Cand cf; cf.gain = KMINSCORE; cf.b = -1; cf.sg = 0; cf.sh = 0; cf.cnt = 0;
  if (two_scans && b <= num_bin - 2) {
    const long long lgi = pg, lhi = ph;
    double gain = 0.0; int lc; bool ok = true;
    if (QUANT) {
      lc = RoundIntD(static_cast<double>(lhi) * cnt_factor);
      const double slh = static_cast<double>(lhi) * hscale;
      const int rc = num_data - lc;
      const long long rgi = tot_g - lgi, rhi = tot_h - lhi;
      const double srh = static_cast<double>(rhi) * hscale;
      ok = !(lc < P.min_data_in_leaf || slh < P.min_sum_hessian) && !(rc < P.min_data_in_leaf) && !(srh < P.min_sum_hessian);
      if (ok) gain = SplitGainD(static_cast<double>(lgi) * gscale, slh, static_cast<double>(rgi) * gscale, srh, P);
    } else {
      lc = pc;
      const double lh = KEPS + static_cast<double>(lhi) * hscale;
      const int rc = num_data - lc;
      const double rh = sum_hessian - lh;
      ok = !(lc < P.min_data_in_leaf || lh < P.min_sum_hessian) && !(rc < P.min_data_in_leaf) && !(rh < P.min_sum_hessian);
      if (ok) { const double lg = static_cast<double>(lgi) * gscale; gain = SplitGainD(lg, lh, sum_gradient - lg, rh, P); }
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
    SplitDev s = EmptySplit(f, fm.real_index);
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

// Leaf selection + Tree::Split bookkeeping, run by one thread block (the last CTA of k_find to finish, so no extra launch):
// picks the leaf with the best split (gain desc, feature asc, leaf asc); stops the tree when no gain is positive;
// otherwise records the new internal node and publishes the split for the partition kernels.
__device__ void SelectAndSplit(LeafDev* __restrict__ leaves, TreeCtrl* __restrict__ ctrl, TreeArraysDev T,
                               const FeatureMetaDev* __restrict__ fmeta, const LearnerParams& P) {
  __shared__ double s_gain[32];
  __shared__ int s_feat[32], s_idx[32];
  const int nl = ctrl->num_leaves;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  double bg = KMINSCORE; int bf = -1, bi = -1;
  for (int l = threadIdx.x; l < nl; l += blockDim.x) {
    const double g = leaves[l].best.gain; const int rf = leaves[l].best.real_feature;
    if (bi < 0 || SplitBetter(g, rf, bg, bf)) { bg = g; bf = rf; bi = l; }   // ascending l: the earlier leaf keeps ties
  }
  auto take = [&](double g, int rf, int i) {
    if (i < 0) return;
    if (bi < 0 || SplitBetter(g, rf, bg, bf) || (!SplitBetter(bg, bf, g, rf) && i < bi)) { bg = g; bf = rf; bi = i; }
  };
  for (int o = 16; o > 0; o >>= 1) {
    const double g = __shfl_xor_sync(0xffffffffu, bg, o);
    const int rf = __shfl_xor_sync(0xffffffffu, bf, o);
    const int i = __shfl_xor_sync(0xffffffffu, bi, o);
    take(g, rf, i);
  }
  __syncthreads();
  if (lane == 0) { s_gain[warp] = bg; s_feat[warp] = bf; s_idx[warp] = bi; }
  __syncthreads();
  if (threadIdx.x != 0) return;
  bg = s_gain[0]; bf = s_feat[0]; bi = s_idx[0];
  for (int w = 1; w < static_cast<int>(blockDim.x >> 5); ++w) take(s_gain[w], s_feat[w], s_idx[w]);
  ctrl->pending = 0;
  const int best_leaf = bi;
  const SplitDev bs = leaves[best_leaf].best;
  if (!(bs.gain > 0.0) || nl >= P.num_leaves) { ctrl->done = 1; return; }
  const int new_node = nl - 1, new_leaf = nl;
  const FeatureMetaDev fm = fmeta[bs.feature];
  const int parent = T.leaf_parent[best_leaf];
  if (parent >= 0) {
    if (T.left_child[parent] == ~best_leaf) T.left_child[parent] = new_node; else T.right_child[parent] = new_node;
  }
  T.split_feature_inner[new_node] = bs.feature;
  T.split_gain[new_node] = static_cast<float>(bs.gain);
  T.left_child[new_node] = ~best_leaf;
  T.right_child[new_node] = ~new_leaf;
  T.leaf_parent[best_leaf] = new_node;
  T.leaf_parent[new_leaf] = new_node;
  T.internal_weight[new_node] = bs.left_sum_h + bs.right_sum_h;
  T.internal_value[new_node] = T.leaf_value[best_leaf];
  T.internal_count[new_node] = bs.left_count + bs.right_count;
  T.leaf_value[best_leaf] = (bs.left_output != bs.left_output) ? 0.0 : bs.left_output;
  T.leaf_weight[best_leaf] = bs.left_sum_h;
  T.leaf_count[best_leaf] = bs.left_count;
  T.leaf_value[new_leaf] = (bs.right_output != bs.right_output) ? 0.0 : bs.right_output;
  T.leaf_weight[new_leaf] = bs.right_sum_h;
  T.leaf_count[new_leaf] = bs.right_count;
  T.leaf_depth[new_leaf] = T.leaf_depth[best_leaf] + 1;
  T.leaf_depth[best_leaf]++;
  signed char dt = 0;
  if (bs.default_left) dt |= 2;
  dt |= static_cast<signed char>((fm.missing_type & 3) << 2);
  T.decision_type[new_node] = dt;
  T.threshold_in_bin[new_node] = bs.threshold;
  ctrl->num_leaves = nl + 1;
  ctrl->cur_leaf = best_leaf; ctrl->new_leaf = new_leaf; ctrl->cur_node = new_node;
  ctrl->cur_feature = bs.feature; ctrl->cur_threshold = bs.threshold; ctrl->cur_default_left = bs.default_left;
  ctrl->cur_nan_bin = (fm.missing_type == kMissingNaN) ? fm.num_bin - 1 : -1;
  ctrl->src_begin = leaves[best_leaf].begin; ctrl->src_count = leaves[best_leaf].count; ctrl->src_buf = leaves[best_leaf].buf;
  ctrl->parent_slot = leaves[best_leaf].hist_slot;
}

// grid = (F, root ? 1 : 2): CTA (f, c) stores child c's histogram of feature f into its pool slot (c = 0: the freshly
// built smaller child; c = 1: the larger child = parent - smaller, K5) and searches it (K6).  The last CTA of each child
// reduces that child's per-feature results to the leaf's best split; the last CTA overall then selects the next leaf to
// split (SelectAndSplit), so one launch covers K5, K6 and the leaf arg-max.
template <bool QUANT>
__global__ void __launch_bounds__(256) k_find(const long long* __restrict__ work, long long* __restrict__ pool,
                                              const FeatureMetaDev* __restrict__ fmeta, LeafDev* __restrict__ leaves,
                                              TreeCtrl* __restrict__ ctrl, SplitDev* __restrict__ feat_best,
                                              unsigned char* __restrict__ splittable, TreeArraysDev T, LearnerParams P, int root) {
  PdlLaunchDependents();
  PdlWait();
  if (ctrl->done) return;
  __shared__ int s_last;
  __shared__ double s_gain[8];
  __shared__ int s_feat[8], s_idx[8];
  if (!root && !ctrl->find) {
    // BeforeFindBestSplit said no: nothing to search, but the next leaf still has to be chosen
    if (blockIdx.x == 0 && blockIdx.y == 0) SelectAndSplit(leaves, ctrl, T, fmeta, P);
    return;
  }
  const int F = P.num_features;
  const int f = blockIdx.x, c = blockIdx.y, b = threadIdx.x;
  const FeatureMetaDev fm = fmeta[f];
  const double gscale = ctrl->grad_scale, hscale = ctrl->hess_scale;
  const size_t slot_stride = static_cast<size_t>(F) * kBins * 2;
  const size_t off = (static_cast<size_t>(f) * kBins + b) * 2;
  const int leaf = root ? 0 : (c == 0 ? ctrl->smaller : ctrl->larger);
  const LeafDev& Lf = leaves[leaf];
  longlong2 v = *reinterpret_cast<const longlong2*>(work + off);
  bool used = true;
  if (!root) {
    used = splittable[static_cast<size_t>(ctrl->cur_leaf) * F + f] != 0;   // the parent's flag (the left child reuses its id)
    if (c == 1) {
      const longlong2 p = *reinterpret_cast<const longlong2*>(pool + static_cast<size_t>(ctrl->parent_slot) * slot_stride + off);
      v.x = p.x - v.x; v.y = p.y - v.y;
    }
  }
  *reinterpret_cast<longlong2*>(pool + static_cast<size_t>(Lf.hist_slot) * slot_stride + off) = v;
  SplitDev* outp = feat_best + static_cast<size_t>(c) * F + f;
  // the parent's flag row is read by both CTAs of feature f and rewritten by one of them: children flags go to a scratch
  // row first and are copied when the child's reduction runs
  unsigned char* spl = splittable + static_cast<size_t>(P.num_leaves + c) * F + f;
  if (!used) {
    if (b == 0) { *outp = EmptySplit(f, fm.real_index); *spl = 0; }
  } else {
    long long tg, th;
    BlockFindBest<QUANT>(v.x, v.y, fm, f, root != 0, Lf.int_g, Lf.int_h, Lf.sum_g, Lf.sum_h, Lf.global_count, P, gscale, hscale,
                         outp, spl, &tg, &th, 0);
    if (root && f == 0 && b == 0) {
      leaves[0].int_g = tg; leaves[0].int_h = th;
      leaves[0].sum_g = static_cast<double>(tg) * gscale; leaves[0].sum_h = static_cast<double>(th) * hscale;
    }
  }
  // ---- last CTA of this child: per-feature results -> the leaf's best split
  __threadfence();
  __syncthreads();
  if (b == 0) s_last = (atomicAdd(&ctrl->find_count[c], 1) == F - 1);
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  {
    const SplitDev* fb = feat_best + static_cast<size_t>(c) * F;
    const unsigned char* sp = splittable + static_cast<size_t>(P.num_leaves + c) * F;
    double bg = KMINSCORE; int bf = -1, bi = -1;
    for (int j = b; j < F; j += blockDim.x) {
      const double g = fb[j].gain; const int rf = fb[j].real_feature;
      if (bi < 0 || SplitBetter(g, rf, bg, bf)) { bg = g; bf = rf; bi = j; }
    }
    const int lane = b & 31, warp = b >> 5;
    for (int o = 16; o > 0; o >>= 1) {
      const double g = __shfl_xor_sync(0xffffffffu, bg, o);
      const int rf = __shfl_xor_sync(0xffffffffu, bf, o);
      const int i = __shfl_xor_sync(0xffffffffu, bi, o);
      if (i >= 0 && (bi < 0 || SplitBetter(g, rf, bg, bf))) { bg = g; bf = rf; bi = i; }
    }
    if (lane == 0) { s_gain[warp] = bg; s_feat[warp] = bf; s_idx[warp] = bi; }
    __syncthreads();
    if (b == 0) {
      bg = s_gain[0]; bf = s_feat[0]; bi = s_idx[0];
      for (int w = 1; w < 8; ++w)
        if (s_idx[w] >= 0 && (bi < 0 || SplitBetter(s_gain[w], s_feat[w], bg, bf))) { bg = s_gain[w]; bf = s_feat[w]; bi = s_idx[w]; }
      leaves[leaf].best = fb[bi];
      ctrl->find_count[c] = 0;
    }
    // children inherit / refresh the per-feature splittable flags
    unsigned char* dstrow = splittable + static_cast<size_t>(leaf) * F;
    // The left child reuses the parent's leaf id, hence its flag row: that row may only be overwritten once every CTA of
    // BOTH children has read it, i.e. by the child reduction that finishes last (below).
    if (root) for (int j = b; j < F; j += blockDim.x) dstrow[j] = sp[j];
    __threadfence();
    __syncthreads();
    if (b == 0) s_last = root ? 1 : (atomicAdd(&ctrl->find_children_done, 1) == 1);
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    if (!root) {
      // all CTAs of both children are done: now it is safe to overwrite the parent's flag row
      for (int cc = 0; cc < 2; ++cc) {
        const int lf2 = cc == 0 ? ctrl->smaller : ctrl->larger;
        const unsigned char* src = splittable + static_cast<size_t>(P.num_leaves + cc) * F;
        unsigned char* dst = splittable + static_cast<size_t>(lf2) * F;
        for (int j = b; j < F; j += blockDim.x) dst[j] = src[j];
      }
      if (b == 0) ctrl->find_children_done = 0;
    }
    __threadfence();
    __syncthreads();
    SelectAndSplit(leaves, ctrl, T, fmeta, P);
  }
}

// ------------------------------------------------------------------------------------------------ K7 stable partition
// A warp owns 32-row words of the leaf's index range: lanes are rows, one ballot gives the word's left/right flags.
// pass 1 (count): flag word per 32 rows + per-block left counts.  pass 2 (scatter): block-local prefix over its words in
// shared memory, then every warp writes its rows to the other index buffer with no further synchronisation; lefts and
// rights each go to consecutive positions, so stores coalesce.  Order is preserved (the partition is stable).
constexpr int kPartWordsPerIter = 2048;     // shared-memory prefix capacity (65536 rows) per block iteration

__device__ __forceinline__ void PartChunk(int n, int* lo_w, int* hi_w) {
  const int words = (n + 31) >> 5;
  const int per = (words + gridDim.x - 1) / gridDim.x;
  *lo_w = min(words, static_cast<int>(blockIdx.x) * per);
  *hi_w = min(words, *lo_w + per);
}

// Body shared by the k_part_count kernel and the persistent tree kernel (any block size that is a multiple of 32, <= 1024).
__device__ __forceinline__ void PartCountBody(const uint8_t* __restrict__ cm, long long n_pad, const int* idx0, const int* idx1,
                                              int es, const TreeCtrl* ctrl, unsigned int* flags, int* block_counts) {
  const int n = ctrl->src_count, begin = ctrl->src_begin, buf = ctrl->src_buf;
  const int thr = ctrl->cur_threshold, nan_bin = ctrl->cur_nan_bin, dl = ctrl->cur_default_left;
  const uint8_t* col = cm + static_cast<size_t>(ctrl->cur_feature) * n_pad;
  const int* idx = buf < 0 ? nullptr : (buf == 0 ? idx0 : idx1);
  int lo_w, hi_w;
  PartChunk(n, &lo_w, &hi_w);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int NW = blockDim.x >> 5;
  int cnt = 0;
  for (int w0 = lo_w + warp; w0 < hi_w; w0 += NW * 4) {
    int rows[4];
    #pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int w = w0 + u * NW;
      const int p = w * 32 + lane;
      rows[u] = (w < hi_w && p < n) ? (idx ? idx[static_cast<size_t>(begin + p) * es] : begin + p) : -1;
    }
    int bins[4];
    #pragma unroll
    for (int u = 0; u < 4; ++u) bins[u] = rows[u] >= 0 ? col[rows[u]] : -1;
    #pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int w = w0 + u * NW;
      const bool left = rows[u] >= 0 && ((bins[u] == nan_bin) ? (dl != 0) : (bins[u] <= thr));
      const unsigned int word = __ballot_sync(0xffffffffu, left);
      if (lane == 0 && w < hi_w) { flags[w] = word; cnt += __popc(word); }
    }
  }
  __shared__ int s_cnt[32];
  __syncthreads();
  if (lane == 0) s_cnt[warp] = cnt;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
    for (int i = 0; i < NW; ++i) t += s_cnt[i];
    block_counts[blockIdx.x] = t;
  }
}

__global__ void __launch_bounds__(kPartThreads) k_part_count(const uint8_t* __restrict__ cm, long long n_pad,
                                                             const int* __restrict__ idx0, const int* __restrict__ idx1, int es,
                                                             const TreeCtrl* __restrict__ ctrl, unsigned int* __restrict__ flags,
                                                             int* __restrict__ block_counts) {
  PdlLaunchDependents();
  PdlWait();
  if (ctrl->done) return;
  PartCountBody(cm, n_pad, idx0, idx1, es, ctrl, flags, block_counts);
}

// Index entries carry the row's gradient pair along with the row id, so a histogram pass over a non-root leaf reads them
// densely instead of gathering 2-8 bytes per row from a sparse set of sectors:
//   quantized (es = 2 ints): {row, (g << 16) + h}        float mode (es = 4 ints): {row, g_fixed, h_fixed, 0}
__device__ __forceinline__ void PartScatterBody(const int* idx0r, const int* idx1r, int* idx0w, int* idx1w, int es,
                                                const uint16_t* __restrict__ gh16, const int2* __restrict__ ghfix, TreeCtrl* ctrl,
                                                LeafDev* leaves, const unsigned int* flags, const int* block_counts,
                                                TreeArraysDev T, long long* work, int work_elems, const LearnerParams& P) {
  // the histogram kernel of this split accumulates into `work` with atomics: clear it here (saves a memset node)
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < work_elems; i += gridDim.x * blockDim.x) work[i] = 0;
  const int n = ctrl->src_count, begin = ctrl->src_begin, buf = ctrl->src_buf;
  const int* idx = buf < 0 ? nullptr : (buf == 0 ? idx0r : idx1r);
  const int other = buf < 0 ? 0 : 1 - buf;
  int* dst = other == 0 ? idx0w : idx1w;
  const int NW = blockDim.x >> 5;
  const int NT = blockDim.x;
  __shared__ int s_red[2][32];
  __shared__ int s_prefix[kPartWordsPerIter];
  __shared__ int s_carry;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int before = 0, total = 0;
  for (int i = threadIdx.x; i < gridDim.x; i += blockDim.x) {
    const int c = block_counts[i];
    total += c;
    if (i < blockIdx.x) before += c;
  }
  for (int o = 16; o > 0; o >>= 1) { before += __shfl_xor_sync(0xffffffffu, before, o); total += __shfl_xor_sync(0xffffffffu, total, o); }
  if (lane == 0) { s_red[0][warp] = before; s_red[1][warp] = total; }
  __syncthreads();
  before = 0; total = 0;
  for (int i = 0; i < NW; ++i) { before += s_red[0][i]; total += s_red[1][i]; }
  int lo_w, hi_w;
  PartChunk(n, &lo_w, &hi_w);
  if (threadIdx.x == 0) s_carry = before;
  __syncthreads();
  for (int base_w = lo_w; base_w < hi_w; base_w += kPartWordsPerIter) {
    const int nw = min(kPartWordsPerIter, hi_w - base_w);
    // exclusive prefix of popcounts over this iteration's words -> s_prefix (lefts before each word)
    for (int t0 = 0; t0 < nw; t0 += NT) {
      const int t = t0 + threadIdx.x;
      int c = t < nw ? __popc(flags[base_w + t]) : 0;
      int incl = c;
      for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += v; }
      if (lane == 31) s_red[0][warp] = incl;
      __syncthreads();
      int wp = 0, tot = 0;
      for (int i = 0; i < NW; ++i) { if (i < warp) wp += s_red[0][i]; tot += s_red[0][i]; }
      if (t < nw) s_prefix[t] = s_carry + wp + incl - c;
      __syncthreads();
      if (threadIdx.x == 0) s_carry += tot;
      __syncthreads();
    }
    // scatter: one word per warp step, 4 words in flight
    for (int w0 = warp; w0 < nw; w0 += NW * 4) {
      int4 ent[4];
      #pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int w = w0 + u * NW;
        const int p = (base_w + w) * 32 + lane;
        ent[u] = make_int4(-1, 0, 0, 0);
        if (w < nw && p < n) {
          if (idx) {
            if (es == 2) { const int2 e = reinterpret_cast<const int2*>(idx)[begin + p]; ent[u].x = e.x; ent[u].y = e.y; }
            else ent[u] = reinterpret_cast<const int4*>(idx)[begin + p];
          } else {           // the root: rows are 0..n-1, gradients come from the per-row arrays
            const int row = begin + p;
            ent[u].x = row;
            if (es == 2) {
              const unsigned int g = gh16[row];
              ent[u].y = static_cast<int>((static_cast<unsigned>(static_cast<int>(static_cast<signed char>(g >> 8))) << 16) + (g & 0xffu));
            } else { const int2 g = ghfix[row]; ent[u].y = g.x; ent[u].z = g.y; }
          }
        }
      }
      #pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int w = w0 + u * NW;
        if (w >= nw) continue;
        const unsigned int word = flags[base_w + w];
        const int p = (base_w + w) * 32 + lane;
        if (ent[u].x >= 0) {
          const int lefts_before = s_prefix[w] + __popc(word & ((1u << lane) - 1u));
          const int pos = ((word >> lane) & 1u) ? begin + lefts_before : begin + total + (p - lefts_before);
          if (es == 2) reinterpret_cast<int2*>(dst)[pos] = make_int2(ent[u].x, ent[u].y);
          else reinterpret_cast<int4*>(dst)[pos] = ent[u];
        }
      }
    }
    __syncthreads();
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    const int left = ctrl->cur_leaf, right = ctrl->new_leaf, node = ctrl->cur_node;
    const SplitDev bs = leaves[left].best;
    const int nl_true = total, nr_true = n - total;
    int lc = bs.left_count, rc = bs.right_count;
    if (!P.dp) {   // serial learner: SplitInfo counts are replaced by the true partition counts
      lc = nl_true; rc = nr_true;
      T.leaf_count[left] = lc; T.leaf_count[right] = rc; T.internal_count[node] = lc + rc;
    }
    LeafDev L = leaves[left], Rr = leaves[right];
    const int fresh_slot = Rr.hist_slot, parent_slot = L.hist_slot, depth = L.depth + 1;
    L.begin = begin; L.count = nl_true; L.buf = other; L.global_count = lc;
    L.int_g = bs.left_int_g; L.int_h = bs.left_int_h; L.sum_g = bs.left_sum_g; L.sum_h = bs.left_sum_h; L.depth = depth;
    Rr.begin = begin + nl_true; Rr.count = nr_true; Rr.buf = other; Rr.global_count = rc;
    Rr.int_g = bs.right_int_g; Rr.int_h = bs.right_int_h; Rr.sum_g = bs.right_sum_g; Rr.sum_h = bs.right_sum_h; Rr.depth = depth;
    L.best.gain = KMINSCORE; L.best.feature = -1; L.best.real_feature = -1;
    Rr.best.gain = KMINSCORE; Rr.best.feature = -1; Rr.best.real_feature = -1;
    const bool left_smaller = lc < rc;
    const int smaller = left_smaller ? left : right, larger = left_smaller ? right : left;
    if (left_smaller) { L.hist_slot = fresh_slot; Rr.hist_slot = parent_slot; }
    // BeforeFindBestSplit
    bool find = true;
    if (P.max_depth > 0 && depth >= P.max_depth) find = false;
    if (rc < P.min_data_in_leaf * 2 && lc < P.min_data_in_leaf * 2) find = false;
    leaves[left] = L; leaves[right] = Rr;
    ctrl->smaller = smaller; ctrl->larger = larger; ctrl->find = find ? 1 : 0; ctrl->pending = find ? 2 : 0;
  }
}


__global__ void __launch_bounds__(kPartThreads) k_part_scatter(const int* __restrict__ idx0r, const int* __restrict__ idx1r,
                                                               int* __restrict__ idx0w, int* __restrict__ idx1w, int es,
                                                               const uint16_t* __restrict__ gh16, const int2* __restrict__ ghfix,
                                                               TreeCtrl* __restrict__ ctrl, LeafDev* __restrict__ leaves,
                                                               const unsigned int* __restrict__ flags,
                                                               const int* __restrict__ block_counts, TreeArraysDev T,
                                                               long long* __restrict__ work, int work_elems, LearnerParams P) {
  PdlLaunchDependents();
  PdlWait();
  if (ctrl->done) return;
  PartScatterBody(idx0r, idx1r, idx0w, idx1w, es, gh16, ghfix, ctrl, leaves, flags, block_counts, T, work, work_elems, P);
}

// ------------------------------------------------------------------------------------------------ K8 score update
// Leaf ranges partition [0, n); ranges are looked up by position.  score[row] += shrunk leaf value.
__global__ void __launch_bounds__(256) k_sort_leaves(const LeafDev* __restrict__ leaves, const TreeCtrl* __restrict__ ctrl,
                                                     int* __restrict__ sorted /*[2*L]: leaf ids then begins*/, int L) {
  const int nl = ctrl->num_leaves;
  for (int i = threadIdx.x; i < nl; i += blockDim.x) {
    const int bi = leaves[i].begin, ci = leaves[i].count;
    int rank = 0;
    for (int j = 0; j < nl; ++j) {
      const int bj = leaves[j].begin, cj = leaves[j].count;
      // order by begin; an empty range sorts before the non-empty range that starts at the same position
      rank += (bj < bi) || (bj == bi && (cj < ci || (cj == ci && j < i)));
    }
    sorted[rank] = i;
    sorted[L + rank] = bi;
  }
}

__global__ void __launch_bounds__(256) k_score_update(double* __restrict__ score, const int* __restrict__ idx0,
                                                      const int* __restrict__ idx1, int es, const LeafDev* __restrict__ leaves,
                                                      const TreeCtrl* __restrict__ ctrl, const int* __restrict__ sorted, int L,
                                                      const double* __restrict__ leaf_value, double lr, int n) {
  __shared__ int s_begin[1024], s_leaf[1024];
  __shared__ double s_val[1024];
  __shared__ int s_buf[1024];
  const int nl = ctrl->num_leaves;
  if (nl <= 1) return;
  for (int i = threadIdx.x; i < nl; i += blockDim.x) {
    const int leaf = sorted[i];
    s_leaf[i] = leaf; s_begin[i] = sorted[L + i]; s_buf[i] = leaves[leaf].buf;
    double v = leaf_value[leaf] * lr;
    const double zt = (double)1e-35f;
    if (v >= -zt && v <= zt) v = 0.0;
    s_val[i] = v;
  }
  __syncthreads();
  for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < n; p += gridDim.x * blockDim.x) {
    int lo = 0, hi = nl - 1;      // last i with s_begin[i] <= p
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (s_begin[mid] <= p) lo = mid; else hi = mid - 1;
    }
    const int row = (s_buf[lo] == 0 ? idx0 : idx1)[static_cast<size_t>(p) * es];
    score[row] += s_val[lo];
  }
}

__global__ void k_tree_init(LeafDev* leaves, TreeCtrl* ctrl, TreeArraysDev T, int L, int n_local, int n_global) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < L) {
    LeafDev z;
    z.begin = 0; z.count = 0; z.buf = -1; z.global_count = 0; z.int_g = z.int_h = 0; z.sum_g = z.sum_h = 0.0;
    z.depth = 0; z.hist_slot = i;
    z.best.gain = KMINSCORE; z.best.feature = -1; z.best.real_feature = -1; z.best.threshold = 0; z.best.default_left = 1;
    z.best.left_count = z.best.right_count = 0; z.best.left_output = z.best.right_output = 0.0;
    z.best.left_sum_g = z.best.left_sum_h = z.best.right_sum_g = z.best.right_sum_h = 0.0;
    z.best.left_int_g = z.best.left_int_h = z.best.right_int_g = z.best.right_int_h = 0;
    if (i == 0) { z.count = n_local; z.global_count = n_global; }
    leaves[i] = z;
    T.leaf_parent[i] = -1; T.leaf_depth[i] = 0; T.leaf_value[i] = 0.0; T.leaf_weight[i] = 0.0; T.leaf_count[i] = 0;
    if (i < L - 1) {
      T.left_child[i] = 0; T.right_child[i] = 0; T.split_feature_inner[i] = 0; T.threshold_in_bin[i] = 0;
      T.internal_count[i] = 0; T.decision_type[i] = 0; T.split_gain[i] = 0.0f; T.internal_value[i] = 0.0; T.internal_weight[i] = 0.0;
    }
  }
  if (i == 0) {
    ctrl->num_leaves = 1; ctrl->done = 0; ctrl->pending = 1; ctrl->find = 1; ctrl->smaller = 0; ctrl->larger = -1;
    ctrl->cur_leaf = 0; ctrl->new_leaf = 0;
    ctrl->find_count[0] = 0; ctrl->find_count[1] = 0; ctrl->find_children_done = 0;
  }
}

}  // namespace lgb
