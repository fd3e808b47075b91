// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_common.h header; parity unpinned).
//
// Objectives, gradient quantisation, histograms, split search, leaf-wise learner (serial and simulated
// data-parallel over S row shards) and the GBDT loop.  Restates upstream
//   src/objective/{binary,regression,rank}_objective.hpp       
//   src/treelearner/gradient_discretizer.cpp                   (A.6.4)
//   src/io/dense_bin.hpp, src/io/dataset.cpp                   (A.5)
//   src/treelearner/feature_histogram.hpp                      (A.7)
//   src/treelearner/{serial,data_parallel}_tree_learner.cpp, data_partition.hpp, leaf_splits.hpp  (A.8)
//   src/boosting/gbdt.cpp, gbdt_model_text.cpp                 (A.9, A.10)
// Reference triggers: LGBM_BoosterCreate (lightgbm/.../TrainUtils.scala:16-29 ->
// booster/LightGBMBooster.scala:241) and LGBM_BoosterUpdateOneIter (TrainUtils.scala:98-135 ->
// LightGBMBooster.scala:355-365); collectives replace the socket learner set up by
// LGBM_NetworkInit (NetworkManager.scala:195-218).
#pragma once
#include <algorithm>
#include <numeric>

#include "orc_dataset.h"
#include "orc_tree.h"

namespace orc {

struct SplitInfo {
  int feature = -1;            // inner feature index
  int real_feature = -1;
  uint32_t threshold = 0;
  data_size_t left_count = 0, right_count = 0;
  double left_output = 0.0, right_output = 0.0;
  double gain = kMinScore;
  double left_sum_gradient = 0.0, left_sum_hessian = 0.0, right_sum_gradient = 0.0, right_sum_hessian = 0.0;
  int64_t left_int_g = 0, left_int_h = 0, right_int_g = 0, right_int_h = 0;
  bool default_left = true;
  int monotone_type = 0;
  std::vector<uint32_t> cat_threshold;      // categorical split: the bins that go left (num_cat_threshold = its size), else empty
  bool categorical = false;
  void Reset() { *this = SplitInfo(); }
  // upstream SplitInfo::operator> : larger gain wins; equal gain -> smaller feature index (-1 counts as +inf)
  bool BetterThan(const SplitInfo& o) const {
    int lf = real_feature == -1 ? INT32_MAX : real_feature;
    int of = o.real_feature == -1 ? INT32_MAX : o.real_feature;
    if (gain != o.gain) return gain > o.gain;
    return lf < of;
  }
};

struct SplitParams {
  double l1, l2, max_delta_step, min_gain_to_split, min_sum_hessian;
  int min_data_in_leaf;
  // categorical split search
  int max_cat_to_onehot = 4, max_cat_threshold = 32, min_data_per_group = 100;
  double cat_smooth = 10.0, cat_l2 = 10.0;
  // monotone constraints, method "basic" (USE_MC is on for every feature once any column is constrained): the leaf's output range
  // and the searched feature's direction
  bool mc = false;
  double cmin = -std::numeric_limits<double>::infinity(), cmax = std::numeric_limits<double>::infinity();
  int mono = 0;
};

inline double ThresholdL1(double s, double l1) {
  double reg = std::max(0.0, std::fabs(s) - l1);
  return Sign(s) * reg;
}
inline double LeafOutput(double g, double h, const SplitParams& p) {
  double ret = (p.l1 > 0.0) ? -ThresholdL1(g, p.l1) / (h + p.l2) : -g / (h + p.l2);
  if (p.max_delta_step > 0.0 && std::fabs(ret) > p.max_delta_step) ret = Sign(ret) * p.max_delta_step;
  return ret;
}
inline double LeafGainGivenOutput(double g, double h, const SplitParams& p, double out) {
  double sg = (p.l1 > 0.0) ? ThresholdL1(g, p.l1) : g;
  return -(2.0 * sg * out + (h + p.l2) * out * out);
}
inline double LeafGain(double g, double h, const SplitParams& p) {
  if (p.max_delta_step <= 0.0) {
    double sg = (p.l1 > 0.0) ? ThresholdL1(g, p.l1) : g;
    return (sg * sg) / (h + p.l2);
  }
  return LeafGainGivenOutput(g, h, p, LeafOutput(g, h, p));
}
// CalculateSplittedLeafOutput<USE_MC>: the output clamped to the leaf's range
inline double LeafOutputC(double g, double h, const SplitParams& p) {
  double ret = LeafOutput(g, h, p);
  if (p.mc) { if (ret < p.cmin) ret = p.cmin; else if (ret > p.cmax) ret = p.cmax; }
  return ret;
}
// GetSplitGains<USE_MC>: with constraints the gain is taken at the clamped outputs, and a split that breaks the feature's direction
// has none
inline double SplitGain(double lg, double lh, double rg, double rh, const SplitParams& p) {
  if (!p.mc) return LeafGain(lg, lh, p) + LeafGain(rg, rh, p);
  const double lo = LeafOutputC(lg, lh, p), ro = LeafOutputC(rg, rh, p);
  if ((p.mono > 0 && lo > ro) || (p.mono < 0 && lo < ro)) return 0;
  return LeafGainGivenOutput(lg, lh, p, lo) + LeafGainGivenOutput(rg, rh, p, ro);
}

// Float-histogram threshold search for one numerical feature (FindBestThresholdSequentially, both directions).
// hist: num_bin (g,h) pairs in logical bin order.  Returns is_splittable.
inline bool FindBestThresholdFloat(const double* hist, const BinMapper& m, double sum_gradient, double sum_hessian_in,
                                   data_size_t num_data, const SplitParams& p, SplitInfo* out) {
  out->default_left = true;
  out->gain = kMinScore;
  const double sum_hessian = sum_hessian_in + 2 * kEpsilon;
  const double gain_shift = LeafGain(sum_gradient, sum_hessian, p);
  const double min_gain_shift = gain_shift + p.min_gain_to_split;
  const int num_bin = m.num_bin;
  const double cnt_factor = num_data / sum_hessian;
  bool splittable = false;
  // skip_default (missing type Zero: FindBestThresholdSequentially<REVERSE, SKIP_DEFAULT_BIN = true, NA_AS_MISSING = false>): the
  // default bin is neither accumulated nor a threshold, so its rows stay on the side the scan does not build.  The forward scan
  // starts at histogram entry 0, which is bin 1 when the most frequent bin is bin 0 (upstream leaves that bin out of the histogram
  // and only the NaN scan reconstructs it).
  auto scan = [&](bool reverse, bool na_as_missing, bool skip_default = false) {
    double best_lg = std::numeric_limits<double>::quiet_NaN(), best_lh = best_lg;
    double best_gain = kMinScore;
    data_size_t best_left_count = 0;
    uint32_t best_threshold = static_cast<uint32_t>(num_bin);
    if (reverse) {
      double rg = 0.0, rh = kEpsilon;
      data_size_t rc = 0;
      for (int b = num_bin - 1 - (na_as_missing ? 1 : 0); b >= 1; --b) {
        if (skip_default && b == static_cast<int>(m.default_bin)) continue;
        double g = hist[2 * b], h = hist[2 * b + 1];
        rg += g; rh += h; rc += RoundInt(h * cnt_factor);
        if (rc < p.min_data_in_leaf || rh < p.min_sum_hessian) continue;
        data_size_t lc = num_data - rc;
        if (lc < p.min_data_in_leaf) break;
        double lh = sum_hessian - rh;
        if (lh < p.min_sum_hessian) break;
        double lg = sum_gradient - rg;
        double gain = SplitGain(lg, lh, rg, rh, p);
        if (gain <= min_gain_shift) continue;
        splittable = true;
        if (gain > best_gain) {
          if (p.mc && p.cmin > p.cmax) continue;
          best_left_count = lc; best_lg = lg; best_lh = lh;
          best_threshold = static_cast<uint32_t>(b - 1);
          best_gain = gain;
        }
      }
    } else {
      double lg = 0.0, lh = kEpsilon;
      data_size_t lc = 0;
      for (int b = (skip_default && m.most_freq_bin == 0) ? 1 : 0; b <= num_bin - 2; ++b) {
        if (skip_default && b == static_cast<int>(m.default_bin)) continue;
        double g = hist[2 * b], h = hist[2 * b + 1];
        lg += g; lh += h; lc += RoundInt(h * cnt_factor);
        if (lc < p.min_data_in_leaf || lh < p.min_sum_hessian) continue;
        data_size_t rc = num_data - lc;
        if (rc < p.min_data_in_leaf) break;
        double rh = sum_hessian - lh;
        if (rh < p.min_sum_hessian) break;
        double rg = sum_gradient - lg;
        double gain = SplitGain(lg, lh, rg, rh, p);
        if (gain <= min_gain_shift) continue;
        splittable = true;
        if (gain > best_gain) {
          if (p.mc && p.cmin > p.cmax) continue;
          best_left_count = lc; best_lg = lg; best_lh = lh;
          best_threshold = static_cast<uint32_t>(b);
          best_gain = gain;
        }
      }
    }
    if (splittable && best_gain > out->gain + min_gain_shift) {
      out->threshold = best_threshold;
      out->left_output = LeafOutputC(best_lg, best_lh, p);
      out->left_count = best_left_count;
      out->left_sum_gradient = best_lg;
      out->left_sum_hessian = best_lh - kEpsilon;
      out->right_output = LeafOutputC(sum_gradient - best_lg, sum_hessian - best_lh, p);
      out->right_count = num_data - best_left_count;
      out->right_sum_gradient = sum_gradient - best_lg;
      out->right_sum_hessian = sum_hessian - best_lh - kEpsilon;
      out->gain = best_gain - min_gain_shift;
      out->default_left = reverse;
    }
  };
  if (num_bin > 2 && m.missing_type == MISSING_ZERO) {
    scan(true, false, true);
    scan(false, false, true);
  } else if (num_bin > 2 && m.missing_type == MISSING_NAN) {
    scan(true, true);
    scan(false, true);
  } else {
    scan(true, false);
    if (m.missing_type == MISSING_NAN) out->default_left = false;
  }
  return splittable;
}

// Integer-histogram threshold search (FindBestThresholdSequentiallyInt): prefix sums are exact integers, gains are
// evaluated on int_sum * scale.
inline bool FindBestThresholdInt(const int64_t* hist, const BinMapper& m, int64_t int_sum_g, int64_t int_sum_h,
                                 double grad_scale, double hess_scale, data_size_t num_data, const SplitParams& p,
                                 SplitInfo* out) {
  out->default_left = true;
  out->gain = kMinScore;
  const double sum_gradient = static_cast<double>(int_sum_g) * grad_scale;
  const double sum_hessian = static_cast<double>(int_sum_h) * hess_scale;
  const double gain_shift = LeafGain(sum_gradient, sum_hessian, p);
  const double min_gain_shift = gain_shift + p.min_gain_to_split;
  const int num_bin = m.num_bin;
  if (int_sum_h <= 0) return false;   // guard (upstream would divide by zero)
  const double cnt_factor = static_cast<double>(num_data) / static_cast<double>(int_sum_h);
  bool splittable = false;
  auto scan = [&](bool reverse, bool na_as_missing, bool skip_default = false) {
    int64_t best_lg = 0, best_lh = 0;
    double best_gain = kMinScore;
    uint32_t best_threshold = static_cast<uint32_t>(num_bin);
    if (reverse) {
      int64_t rg = 0, rh = 0;
      for (int b = num_bin - 1 - (na_as_missing ? 1 : 0); b >= 1; --b) {
        if (skip_default && b == static_cast<int>(m.default_bin)) continue;
        rg += hist[2 * b]; rh += hist[2 * b + 1];
        data_size_t rc = RoundInt(static_cast<double>(rh) * cnt_factor);
        double srh = static_cast<double>(rh) * hess_scale;
        if (rc < p.min_data_in_leaf || srh < p.min_sum_hessian) continue;
        data_size_t lc = num_data - rc;
        if (lc < p.min_data_in_leaf) break;
        int64_t lg = int_sum_g - rg, lh = int_sum_h - rh;
        double slh = static_cast<double>(lh) * hess_scale;
        if (slh < p.min_sum_hessian) break;
        double srg = static_cast<double>(rg) * grad_scale, slg = static_cast<double>(lg) * grad_scale;
        double gain = SplitGain(slg, slh, srg, srh, p);
        if (gain <= min_gain_shift) continue;
        splittable = true;
        if (gain > best_gain && !(p.mc && p.cmin > p.cmax)) { best_lg = lg; best_lh = lh; best_threshold = static_cast<uint32_t>(b - 1); best_gain = gain; }
      }
    } else {
      int64_t lg = 0, lh = 0;
      for (int b = (skip_default && m.most_freq_bin == 0) ? 1 : 0; b <= num_bin - 2; ++b) {
        if (skip_default && b == static_cast<int>(m.default_bin)) continue;
        lg += hist[2 * b]; lh += hist[2 * b + 1];
        data_size_t lc = RoundInt(static_cast<double>(lh) * cnt_factor);
        double slh = static_cast<double>(lh) * hess_scale;
        if (lc < p.min_data_in_leaf || slh < p.min_sum_hessian) continue;
        data_size_t rc = num_data - lc;
        if (rc < p.min_data_in_leaf) break;
        int64_t rg = int_sum_g - lg, rh = int_sum_h - lh;
        double srh = static_cast<double>(rh) * hess_scale;
        if (srh < p.min_sum_hessian) break;
        double srg = static_cast<double>(rg) * grad_scale, slg = static_cast<double>(lg) * grad_scale;
        double gain = SplitGain(slg, slh, srg, srh, p);
        if (gain <= min_gain_shift) continue;
        splittable = true;
        if (gain > best_gain && !(p.mc && p.cmin > p.cmax)) { best_lg = lg; best_lh = lh; best_threshold = static_cast<uint32_t>(b); best_gain = gain; }
      }
    }
    if (splittable && best_gain > out->gain + min_gain_shift) {
      int64_t rg = int_sum_g - best_lg, rh = int_sum_h - best_lh;
      double blg = static_cast<double>(best_lg) * grad_scale, blh = static_cast<double>(best_lh) * hess_scale;
      double brg = static_cast<double>(rg) * grad_scale, brh = static_cast<double>(rh) * hess_scale;
      out->threshold = best_threshold;
      out->left_output = LeafOutputC(blg, blh, p);
      out->left_count = RoundInt(static_cast<double>(best_lh) * cnt_factor);
      out->left_sum_gradient = blg; out->left_sum_hessian = blh;
      out->left_int_g = best_lg; out->left_int_h = best_lh;
      out->right_output = LeafOutputC(brg, brh, p);
      out->right_count = RoundInt(static_cast<double>(rh) * cnt_factor);
      out->right_sum_gradient = brg; out->right_sum_hessian = brh;
      out->right_int_g = rg; out->right_int_h = rh;
      out->gain = best_gain - min_gain_shift;
      out->default_left = reverse;
    }
  };
  if (num_bin > 2 && m.missing_type == MISSING_ZERO) {
    scan(true, false, true);
    scan(false, false, true);
  } else if (num_bin > 2 && m.missing_type == MISSING_NAN) {
    scan(true, true);
    scan(false, true);
  } else {
    scan(true, false);
    if (m.missing_type == MISSING_NAN) out->default_left = false;
  }
  return splittable;
}

// Categorical split search (FindBestThresholdCategoricalInner / ...IntInner).  Bin 0 (NaN, negative, rare and unseen categories) is
// never a candidate and always goes right.  Few bins (num_bin <= max_cat_to_onehot): every category alone against the rest.
// Otherwise the bins with at least cat_smooth rows are sorted by sum_gradient / (sum_hessian + cat_smooth) and prefixes of that
// order are tried from both ends (at most max_cat_threshold categories, groups of at least min_data_per_group rows), with cat_l2
// added to lambda_l2.  QUANT: hist holds integers and every sum is int_sum * scale, without the kEpsilon terms of the float form.
template <bool QUANT, typename HistT>
inline bool FindBestThresholdCategorical(const HistT* hist, const BinMapper& m, double leaf_sum_g, double leaf_sum_h, int64_t int_sum_g,
                                         int64_t int_sum_h, double grad_scale, double hess_scale, data_size_t num_data,
                                         const SplitParams& p_in, SplitInfo* out) {
  out->default_left = false;
  out->gain = kMinScore;
  out->categorical = true;
  SplitParams p = p_in;
  double sum_gradient, sum_hessian, cnt_factor;
  if (QUANT) {
    if (int_sum_h <= 0) return false;
    sum_gradient = static_cast<double>(int_sum_g) * grad_scale;
    sum_hessian = static_cast<double>(int_sum_h) * hess_scale;
    cnt_factor = static_cast<double>(num_data) / static_cast<double>(int_sum_h);
  } else {
    sum_gradient = leaf_sum_g;
    sum_hessian = leaf_sum_h + 2 * kEpsilon;
    cnt_factor = num_data / sum_hessian;
  }
  const double eps = QUANT ? 0.0 : kEpsilon;
  const double gain_shift = LeafGain(sum_gradient, sum_hessian, p);
  const double min_gain_shift = gain_shift + p.min_gain_to_split;
  const int num_bin = m.num_bin;
  auto grad_of = [&](int b) { return QUANT ? static_cast<double>(hist[2 * b]) * grad_scale : static_cast<double>(hist[2 * b]); };
  auto hess_of = [&](int b) { return QUANT ? static_cast<double>(hist[2 * b + 1]) * hess_scale : static_cast<double>(hist[2 * b + 1]); };
  auto cnt_of = [&](int b) { return QUANT ? RoundInt(static_cast<double>(hist[2 * b + 1]) * cnt_factor) : RoundInt(hess_of(b) * cnt_factor); };
  bool splittable = false;
  double best_gain = kMinScore;
  data_size_t best_left_count = 0;
  double best_lg = 0, best_lh = 0;
  int64_t best_int_lg = 0, best_int_lh = 0;
  int best_threshold = -1, best_dir = 1;
  const bool use_onehot = num_bin <= p.max_cat_to_onehot;
  std::vector<int> sorted_idx;
  int used_bin = -1;
  if (use_onehot) {
    for (int t = 1; t < num_bin; ++t) {
      const double grad = grad_of(t), hess = hess_of(t);
      const data_size_t cnt = cnt_of(t);
      if (cnt < p.min_data_in_leaf || hess < p.min_sum_hessian) continue;
      const data_size_t other_count = num_data - cnt;
      if (other_count < p.min_data_in_leaf) continue;
      const double sum_other_hessian = QUANT ? static_cast<double>(int_sum_h - static_cast<int64_t>(hist[2 * t + 1])) * hess_scale
                                             : sum_hessian - hess - kEpsilon;
      if (sum_other_hessian < p.min_sum_hessian) continue;
      const double sum_other_gradient = QUANT ? static_cast<double>(int_sum_g - static_cast<int64_t>(hist[2 * t])) * grad_scale
                                              : sum_gradient - grad;
      const double gain = SplitGain(sum_other_gradient, sum_other_hessian, grad, hess + eps, p);
      if (gain <= min_gain_shift) continue;
      splittable = true;
      if (gain > best_gain) {
        best_threshold = t; best_lg = grad; best_lh = hess + eps; best_left_count = cnt; best_gain = gain;
        if (QUANT) { best_int_lg = static_cast<int64_t>(hist[2 * t]); best_int_lh = static_cast<int64_t>(hist[2 * t + 1]); }
      }
    }
  } else {
    for (int i = 1; i < num_bin; ++i)
      if (cnt_of(i) >= p.cat_smooth) sorted_idx.push_back(i);
    used_bin = static_cast<int>(sorted_idx.size());
    p.l2 += p.cat_l2;
    auto ctr = [&](int b) { return grad_of(b) / (hess_of(b) + p.cat_smooth); };
    std::stable_sort(sorted_idx.begin(), sorted_idx.end(), [&](int i, int j) { return ctr(i) < ctr(j); });
    const int max_num_cat = std::min(p.max_cat_threshold, (used_bin + 1) / 2);
    for (int pass = 0; pass < 2; ++pass) {
      const int dir = pass == 0 ? 1 : -1;
      int start_pos = pass == 0 ? 0 : used_bin - 1;
      data_size_t cnt_cur_group = 0, left_count = 0;
      double sum_left_gradient = 0.0, sum_left_hessian = kEpsilon;
      int64_t int_lg = 0, int_lh = 0;
      for (int i = 0; i < used_bin && i < max_num_cat; ++i) {
        const int t = sorted_idx[start_pos];
        start_pos += dir;
        const data_size_t cnt = cnt_of(t);
        if (QUANT) {
          int_lg += static_cast<int64_t>(hist[2 * t]); int_lh += static_cast<int64_t>(hist[2 * t + 1]);
          sum_left_gradient = static_cast<double>(int_lg) * grad_scale;
          sum_left_hessian = static_cast<double>(int_lh) * hess_scale;
        } else {
          sum_left_gradient += grad_of(t);
          sum_left_hessian += hess_of(t);
        }
        left_count += cnt;
        cnt_cur_group += cnt;
        if (left_count < p.min_data_in_leaf || sum_left_hessian < p.min_sum_hessian) continue;
        const data_size_t right_count = num_data - left_count;
        if (right_count < p.min_data_in_leaf || right_count < p.min_data_per_group) break;
        const double sum_right_hessian = QUANT ? static_cast<double>(int_sum_h - int_lh) * hess_scale : sum_hessian - sum_left_hessian;
        if (sum_right_hessian < p.min_sum_hessian) break;
        if (cnt_cur_group < p.min_data_per_group) continue;
        cnt_cur_group = 0;
        const double sum_right_gradient = QUANT ? static_cast<double>(int_sum_g - int_lg) * grad_scale : sum_gradient - sum_left_gradient;
        const double gain = SplitGain(sum_left_gradient, sum_left_hessian, sum_right_gradient, sum_right_hessian, p);
        if (gain <= min_gain_shift) continue;
        splittable = true;
        if (gain > best_gain) {
          best_left_count = left_count; best_lg = sum_left_gradient; best_lh = sum_left_hessian;
          best_int_lg = int_lg; best_int_lh = int_lh;
          best_threshold = i; best_gain = gain; best_dir = dir;
        }
      }
    }
  }
  if (splittable) {
    out->left_count = best_left_count;
    out->right_count = num_data - best_left_count;
    if (QUANT) {
      const int64_t rg = int_sum_g - best_int_lg, rh = int_sum_h - best_int_lh;
      const double brg = static_cast<double>(rg) * grad_scale, brh = static_cast<double>(rh) * hess_scale;
      out->left_output = LeafOutputC(best_lg, best_lh, p);
      out->left_sum_gradient = best_lg; out->left_sum_hessian = best_lh;
      out->left_int_g = best_int_lg; out->left_int_h = best_int_lh;
      out->right_output = LeafOutputC(brg, brh, p);
      out->right_sum_gradient = brg; out->right_sum_hessian = brh;
      out->right_int_g = rg; out->right_int_h = rh;
    } else {
      out->left_output = LeafOutputC(best_lg, best_lh, p);
      out->left_sum_gradient = best_lg; out->left_sum_hessian = best_lh - kEpsilon;
      out->right_output = LeafOutputC(sum_gradient - best_lg, sum_hessian - best_lh, p);
      out->right_sum_gradient = sum_gradient - best_lg; out->right_sum_hessian = sum_hessian - best_lh - kEpsilon;
    }
    out->gain = best_gain - min_gain_shift;
    if (use_onehot) {
      out->cat_threshold.assign(1, static_cast<uint32_t>(best_threshold));
    } else {
      out->cat_threshold.resize(best_threshold + 1);
      for (int i = 0; i <= best_threshold; ++i)
        out->cat_threshold[i] = static_cast<uint32_t>(best_dir == 1 ? sorted_idx[i] : sorted_idx[used_bin - 1 - i]);
    }
  }
  return splittable;
}

// PercentileFun / WeightedPercentileFun (upstream src/objective/regression_objective.hpp), on a copy of the data.  alpha_f: the
// quantile objective keeps alpha as a float, so (1.0f - alpha) * cnt is float arithmetic there; L1 uses the double 0.5.
template <typename T>
inline T PercentileOf(std::vector<T> data, double alpha, bool alpha_is_float) {
  const int cnt = static_cast<int>(data.size());
  if (cnt <= 1) return data[0];
  const double float_pos = alpha_is_float ? static_cast<double>((1.0f - static_cast<float>(alpha)) * cnt) : (1.0f - alpha) * cnt;
  const int pos = static_cast<int>(float_pos);
  std::sort(data.begin(), data.end(), std::greater<T>());      // descending: [pos - 1] is the pos-th largest
  if (pos < 1) return data[0];
  if (pos >= cnt) return data[cnt - 1];
  const double bias = float_pos - pos;
  const T v1 = data[pos - 1], v2 = data[pos];
  return static_cast<T>(v1 - (v1 - v2) * bias);
}
template <typename T1, typename T2>
inline T1 WeightedPercentileOf(const std::vector<T1>& data, const std::vector<T2>& weight, double alpha, bool alpha_is_float) {
  const int cnt = static_cast<int>(data.size());
  if (cnt <= 1) return data[0];
  std::vector<int> idx(cnt);
  for (int i = 0; i < cnt; ++i) idx[i] = i;
  std::stable_sort(idx.begin(), idx.end(), [&](int a, int b) { return data[a] < data[b]; });
  std::vector<T2> cdf(cnt);
  cdf[0] = weight[idx[0]];
  for (int i = 1; i < cnt; ++i) cdf[i] = cdf[i - 1] + weight[idx[i]];
  const T2 threshold = alpha_is_float ? static_cast<T2>(cdf[cnt - 1] * static_cast<float>(alpha)) : static_cast<T2>(cdf[cnt - 1] * alpha);
  size_t pos = std::upper_bound(cdf.begin(), cdf.end(), threshold) - cdf.begin();
  pos = std::min(pos, static_cast<size_t>(cnt - 1));
  if (pos == 0 || pos == static_cast<size_t>(cnt - 1)) return data[idx[pos]];
  const T1 v1 = data[idx[pos - 1]], v2 = data[idx[pos]];
  if (cdf[pos + 1] - cdf[pos] >= 1.0f) return static_cast<T1>((threshold - cdf[pos]) / (cdf[pos + 1] - cdf[pos]) * (v2 - v1) + v1);
  return static_cast<T1>(v2);
}

struct Shard {
  const Dataset* ds = nullptr;
  int64_t row_offset = 0;                 // global row id of local row 0 (for the stochastic-rounding stream)
  std::vector<double> score;              // scores of the class being trained (multiclass: swapped with score_k[k] around its tree)
  std::vector<score_t> grad, hess;        // gradients of the tree being trained
  std::vector<std::vector<double>> score_k;           // multiclass: [class][row]
  std::vector<std::vector<score_t>> grad_k, hess_k;   // multiclass: [class][row], filled once per iteration
  std::vector<int8_t> qg, qh;             // quantized gradient / hessian
  std::vector<data_size_t> indices, tmp_left, tmp_right;
  std::vector<data_size_t> leaf_begin, leaf_count;
  // bagging (upstream src/boosting/bagging.hpp): in-bag rows ascending in [0, bag_cnt), the others behind them
  std::vector<data_size_t> bag_indices;
  data_size_t bag_cnt = 0;
  std::vector<Random> bag_rands;          // one stream per block of 1024 rows, seeded bagging_seed + block
};

class Booster {
 public:
  Config cfg;
  std::vector<Shard> shards;
  int S = 1, F = 0, B = 1;
  bool dp = false;                         // data-parallel semantics (global estimated counts) when S > 1
  std::vector<std::unique_ptr<Tree>> models;
  int iter = 0;
  // objective state
  enum Obj { OBJ_BINARY, OBJ_L2, OBJ_LAMBDARANK, OBJ_HUBER, OBJ_FAIR, OBJ_POISSON, OBJ_GAMMA, OBJ_TWEEDIE, OBJ_CUSTOM, OBJ_MULTICLASS, OBJ_L1, OBJ_QUANTILE, OBJ_MAPE, OBJ_MULTICLASSOVA, OBJ_XENDCG, OBJ_XENTROPY } obj = OBJ_L2;
  std::vector<std::vector<Random>> xe_rands;      // rank_xendcg: per shard, one stream per query (seeded objective_seed + query)
  std::vector<int> mono_inner;             // monotone direction per inner feature (empty: no constraints at all)
  bool IsMulti() const { return obj == OBJ_MULTICLASS || obj == OBJ_MULTICLASSOVA || (obj == OBJ_CUSTOM && K > 1); }
  std::vector<double> ova_label_weights, ova_suml;      // multiclassova: per class [neg, pos] label weights and the weighted positives
  double ova_sumw = 0.0;
  bool RenewsLeaves() const { return obj == OBJ_L1 || obj == OBJ_QUANTILE || obj == OBJ_MAPE; }      // IsRenewTreeOutput
  std::vector<std::vector<label_t>> label_weight;      // mape: per shard, 1 / max(1, |label|) (* weight)
  enum Boost { BOOST_GBDT, BOOST_RF, BOOST_DART, BOOST_GOSS } boost = BOOST_GBDT;
  double shrinkage_rate = 0.1;             // learning_rate, except under DART, which rescales it by the number of dropped trees
  std::vector<double> rf_init_scores;      // RF: the automatic initial scores (never added to the scores; every tree carries them)
  Random drop_rand;                        // DART: random_for_drop_
  std::vector<double> tree_weight;         // DART: one weight per iteration
  double sum_weight = 0.0;
  std::vector<int> drop_index;
  int K = 1;                               // trees per iteration (num_class for multiclass)
  int cur_class = 0;
  std::vector<double> class_init_probs;    // multiclass: (weighted) share of every label over all machines
  std::vector<char> class_need_train;
  bool IsPointRegression() const { return obj == OBJ_L2 || (obj >= OBJ_HUBER && obj <= OBJ_TWEEDIE) || RenewsLeaves(); }
  bool LogLink() const { return obj == OBJ_POISSON || obj == OBJ_GAMMA || obj == OBJ_TWEEDIE; }
  bool need_train = true;
  double label_weights[2] = {1.0, 1.0};
  bool has_weights = false;
  std::vector<double> label_gain, discount, sigmoid_table;
  std::vector<std::vector<double>> inverse_max_dcgs;   // per shard per query
  double min_sigmoid_input = -50, max_sigmoid_input = 50, sigmoid_table_idx_factor = 1;
  // learner state
  SplitParams sp;
  double grad_scale = 1.0, hess_scale = 1.0;
  struct Leaf {
    double sum_g = 0, sum_h = 0; int64_t int_g = 0, int_h = 0;
    data_size_t count = 0;               // count used by the split finder (serial: true; dp: global estimate)
    int depth = 0;
    double cmin = -std::numeric_limits<double>::infinity(), cmax = std::numeric_limits<double>::infinity();      // BasicLeafConstraints entry
    SplitInfo best;
    std::vector<char> splittable;
  };
  std::vector<Leaf> leaves;
  std::vector<std::vector<double>> hist_f;   // per leaf: F*B*2 doubles
  std::vector<std::vector<int64_t>> hist_i;  // per leaf: F*B*2 int64
  int64_t total_rows = 0;
  // column / row sampling
  Random feature_rand;
  std::vector<char> tree_mask;             // this tree's column sample (all ones when feature_fraction == 1)
  bool bagging = false, balanced_bagging = false;

  Booster(const std::vector<const Dataset*>& ds, const char* params) {
    cfg = ds[0]->config;     // dataset-time params (max_bin ...) first, booster string overrides
    cfg.Set(params ? params : "");
    S = static_cast<int>(ds.size());
    dp = S > 1;
    F = ds[0]->num_features();
    B = ds[0]->max_num_bin;
    shards.resize(S);
    int64_t off = 0;
    for (int s = 0; s < S; ++s) {
      shards[s].ds = ds[s];
      shards[s].row_offset = off;
      off += ds[s]->num_data;
      int64_t n = ds[s]->num_data;
      shards[s].score.assign(n, 0.0);
      shards[s].grad.assign(n, 0.0f);
      shards[s].hess.assign(n, 0.0f);
      if (cfg.use_quantized_grad) { shards[s].qg.assign(n, 0); shards[s].qh.assign(n, 0); }
      shards[s].indices.resize(n); shards[s].tmp_left.resize(n); shards[s].tmp_right.resize(n);
      shards[s].leaf_begin.assign(cfg.num_leaves, 0); shards[s].leaf_count.assign(cfg.num_leaves, 0);
      if (ds[s]->has_init_score) for (int64_t i = 0; i < n; ++i) shards[s].score[i] = ds[s]->init_score[i];
    }
    total_rows = off;
    has_weights = ds[0]->has_weight;
    if (std::any_of(cfg.monotone_constraints.begin(), cfg.monotone_constraints.end(), [](int v) { return v != 0; })) {
      if (cfg.monotone_constraints_method != "basic") throw std::runtime_error("oracle: only monotone_constraints_method=basic is restated");
      if (static_cast<int>(cfg.monotone_constraints.size()) != ds[0]->num_total_features) throw std::runtime_error("oracle: monotone_constraints needs one value per column");
      for (int f = 0; f < F; ++f) mono_inner.push_back(cfg.monotone_constraints[ds[0]->used_features[f]]);
    }
    sp = {cfg.lambda_l1, cfg.lambda_l2, cfg.max_delta_step, cfg.min_gain_to_split, cfg.min_sum_hessian_in_leaf, cfg.min_data_in_leaf,
          cfg.max_cat_to_onehot, cfg.max_cat_threshold, cfg.min_data_per_group, cfg.cat_smooth, cfg.cat_l2};
    if (cfg.boosting == "gbdt") boost = BOOST_GBDT;
    else if (cfg.boosting == "rf") boost = BOOST_RF;
    else if (cfg.boosting == "dart") boost = BOOST_DART;
    else if (cfg.boosting == "goss") boost = BOOST_GOSS;
    else throw std::runtime_error("oracle: boosting=" + cfg.boosting + " is not restated");
    shrinkage_rate = boost == BOOST_RF ? 1.0 : cfg.learning_rate;
    drop_rand = Random(cfg.drop_seed);
    feature_rand = Random(cfg.feature_fraction_seed);
    tree_mask.assign(F, 1);
    // ResetSampleConfig: balanced bagging needs a binary objective (NumPositiveData) with positive rows
    balanced_bagging = cfg.bagging_freq > 0 && (cfg.pos_bagging_fraction < 1.0 || cfg.neg_bagging_fraction < 1.0) && cfg.objective == "binary";
    if (balanced_bagging) {
      bool any_pos = false;
      for (auto d : ds) for (int64_t i = 0; i < d->num_data && !any_pos; ++i) any_pos = d->label[i] > 0;
      balanced_bagging = any_pos;
    }
    bagging = (cfg.bagging_freq > 0 && (cfg.bagging_fraction < 1.0 || balanced_bagging)) || boost == BOOST_GOSS;
    if (boost == BOOST_RF && !(bagging || cfg.feature_fraction < 1.0))
      throw std::runtime_error("oracle: boosting=rf needs bagging (bagging_freq > 0, bagging_fraction < 1) or feature_fraction < 1");
    for (auto& sh : shards) {
      const data_size_t n = static_cast<data_size_t>(sh.ds->num_data);
      sh.bag_cnt = n;
      if (bagging) {       // ResetSampleConfig: every machine seeds its blocks alike (bagging is local to a machine)
        sh.bag_indices.resize(n);
        for (data_size_t b = 0; b < (n + 1023) / 1024; ++b) sh.bag_rands.emplace_back(cfg.bagging_seed + static_cast<int>(b));
      }
    }
    InitObjective();
  }

  // BaggingSampleStrategy::Bagging + BaggingHelper: at iterations that are multiples of bagging_freq every row draws once from
  // the stream of its 1024-row block and is in the bag when the draw is below bagging_fraction.  (The subset copy upstream makes
  // for small bags changes no tree; the discretizer then sees the bag only — this restatement always uses all rows for the scales.)
  void Bagging() {
    if (boost == BOOST_GOSS) { GossSample(); return; }
    if (!bagging || iter % cfg.bagging_freq != 0) return;
    for (auto& sh : shards) {
      const data_size_t n = static_cast<data_size_t>(sh.ds->num_data);
      data_size_t left = 0, right = n;
      for (data_size_t i = 0; i < n; ++i) {
        // BalancedBaggingHelper: positives and negatives of a binary problem are kept with their own fractions
        const double fraction = !balanced_bagging ? cfg.bagging_fraction : sh.ds->label[i] > 0 ? cfg.pos_bagging_fraction : cfg.neg_bagging_fraction;
        if (sh.bag_rands[i / 1024].NextFloat() < fraction) sh.bag_indices[left++] = i;
        else sh.bag_indices[--right] = i;
      }
      sh.bag_cnt = left;
    }
  }

  // GOSS (upstream src/boosting/goss.hpp, BaggingHelper): no sampling during the first 1 / learning_rate iterations; afterwards, per
  // block of rows, the top_rate share with the largest sum over classes of |g * h| is kept and other_rate of the rest is drawn
  // sequentially from the block's LCG stream, its gradients scaled by (cnt - top_k) / other_k.  Upstream cuts the rows into blocks
  // by its OpenMP thread count (at least 1024 rows each); this restatement uses one block per 1024 rows.
  void GossSample() {
    for (auto& sh : shards) {
      const data_size_t n = static_cast<data_size_t>(sh.ds->num_data);
      if (iter < static_cast<int>(1.0f / cfg.learning_rate)) {
        for (data_size_t i = 0; i < n; ++i) sh.bag_indices[i] = i;
        sh.bag_cnt = n;
        continue;
      }
      std::vector<std::pair<std::vector<data_size_t>, std::vector<data_size_t>>> parts((n + 1023) / 1024);
#pragma omp parallel for schedule(dynamic)
      for (data_size_t b = 0; b < (n + 1023) / 1024; ++b) {
        const data_size_t start = b * 1024, cnt = std::min<data_size_t>(1024, n - start);
        std::vector<score_t> tmp(cnt, 0.0f);
        auto mag = [&](data_size_t i) {
          score_t a = 0.0f;
          if (K == 1) a += std::fabs(sh.grad[start + i] * sh.hess[start + i]);
          else for (int k = 0; k < K; ++k) a += std::fabs(sh.grad_k[k][start + i] * sh.hess_k[k][start + i]);
          return a;
        };
        for (data_size_t i = 0; i < cnt; ++i) tmp[i] = mag(i);
        data_size_t top_k = static_cast<data_size_t>(cnt * cfg.top_rate);
        const data_size_t other_k = static_cast<data_size_t>(cnt * cfg.other_rate);
        top_k = std::max<data_size_t>(1, top_k);
        std::vector<score_t> sorted(tmp);
        std::nth_element(sorted.begin(), sorted.begin() + (top_k - 1), sorted.end(), std::greater<score_t>());      // ArgMaxAtK
        const score_t threshold = sorted[top_k - 1];
        const score_t multiply = static_cast<score_t>(cnt - top_k) / other_k;
        data_size_t big = 0;
        auto& left = parts[b].first; auto& right = parts[b].second;
        for (data_size_t i = 0; i < cnt; ++i) {
          if (tmp[i] >= threshold) { left.push_back(start + i); ++big; continue; }
          const data_size_t sampled = static_cast<data_size_t>(left.size()) - big;
          const data_size_t rest_need = other_k - sampled;
          const data_size_t rest_all = (cnt - i) - (top_k - big);
          const double prob = rest_need / static_cast<double>(rest_all);
          if (sh.bag_rands[b].NextFloat() < prob) {
            left.push_back(start + i);
            if (K == 1) { sh.grad[start + i] *= multiply; sh.hess[start + i] *= multiply; }
            else for (int k = 0; k < K; ++k) { sh.grad_k[k][start + i] *= multiply; sh.hess_k[k][start + i] *= multiply; }
          } else {
            right.push_back(start + i);
          }
        }
      }
      data_size_t l = 0, r = n;
      for (auto& p : parts) {      // in-bag rows ascending in front, the others behind them
        for (data_size_t v : p.first) sh.bag_indices[l++] = v;
        for (data_size_t v : p.second) sh.bag_indices[--r] = v;
      }
      sh.bag_cnt = l;
    }
  }

  // SerialTreeLearner::RenewTreeOutput with the objective's RenewTreeOutput (l1 / quantile / mape): every leaf's output becomes the
  // (weighted) percentile of label - score over the leaf's rows on each machine; machines that hold rows of the leaf are averaged.
  // const_pred: a random forest re-fits against label - initial score (RF::TrainOneIter's residual_getter), boosting against label - score
  void RenewTreeOutput(Tree* tree, const double* const_pred = nullptr) {
    const double a = obj == OBJ_QUANTILE ? cfg.alpha : 0.5;
    const bool af = obj == OBJ_QUANTILE;
    for (int l = 0; l < tree->num_leaves; ++l) {
      double acc = 0.0;
      int nworkers = 0;
      for (size_t s = 0; s < shards.size(); ++s) {
        Shard& sh = shards[s];
        const data_size_t b = sh.leaf_begin[l], n = sh.leaf_count[l];
        if (n <= 0) continue;
        std::vector<double> res(n);
        std::vector<label_t> w(n);
        for (data_size_t i = 0; i < n; ++i) {
          const data_size_t r = sh.indices[b + i];
          res[i] = static_cast<double>(sh.ds->label[r]) - (const_pred != nullptr ? *const_pred : sh.score[r]);
          if (obj == OBJ_MAPE) w[i] = label_weight[s][r]; else if (has_weights) w[i] = sh.ds->weight[r];
        }
        double out;
        if (obj == OBJ_MAPE || has_weights) out = WeightedPercentileOf<double, label_t>(res, w, a, af);
        else out = PercentileOf<double>(res, a, af);
        acc += out;
        ++nworkers;
      }
      tree->leaf_value[l] = shards.size() > 1 ? acc / nworkers : acc;
    }
  }

  // ScoreUpdater::AddScore(tree, class): every row walks the tree over its bins
  void AddTreeToScores(const Tree& t, int k) {
    for (auto& sh : shards) {
      std::vector<double>& sc = IsMulti() ? sh.score_k[k] : sh.score;
      const int64_t n = sh.ds->num_data;
#pragma omp parallel for schedule(static)
      for (int64_t i = 0; i < n; ++i) sc[i] += t.leaf_value[LeafByBins(t, sh.ds, static_cast<data_size_t>(i))];
    }
  }

  // DART::DroppingTrees (upstream src/boosting/dart.hpp): picks the iterations to drop, takes their trees out of the training
  // scores and sets this iteration's shrinkage
  void DroppingTrees() {
    drop_index.clear();
    const bool is_skip = drop_rand.NextFloat() < cfg.skip_drop;
    if (!is_skip) {
      double drop_rate = cfg.drop_rate;
      if (!cfg.uniform_drop) {
        const double inv_average_weight = static_cast<double>(tree_weight.size()) / sum_weight;
        if (cfg.max_drop > 0) drop_rate = std::min(drop_rate, cfg.max_drop * inv_average_weight / sum_weight);
        for (int i = 0; i < iter; ++i) {
          if (drop_rand.NextFloat() < drop_rate * tree_weight[i] * inv_average_weight) {
            drop_index.push_back(i);
            if (drop_index.size() >= static_cast<size_t>(cfg.max_drop)) break;
          }
        }
      } else {
        if (cfg.max_drop > 0) drop_rate = std::min(drop_rate, cfg.max_drop / static_cast<double>(iter));
        for (int i = 0; i < iter; ++i) {
          if (drop_rand.NextFloat() < drop_rate) {
            drop_index.push_back(i);
            if (drop_index.size() >= static_cast<size_t>(cfg.max_drop)) break;
          }
        }
      }
    }
    for (int i : drop_index) {
      for (int k = 0; k < K; ++k) {
        Tree& t = *models[static_cast<size_t>(i) * K + k];
        t.Shrinkage(-1.0);
        AddTreeToScores(t, k);
      }
    }
    if (!cfg.xgboost_dart_mode) shrinkage_rate = cfg.learning_rate / (1.0f + static_cast<double>(drop_index.size()));
    else if (drop_index.empty()) shrinkage_rate = cfg.learning_rate;
    else shrinkage_rate = cfg.learning_rate / (cfg.learning_rate + static_cast<double>(drop_index.size()));
  }

  // DART::Normalize: the dropped trees come back scaled by k / (k + 1) (xgboost mode: k / (k + learning_rate))
  void Normalize() {
    const double k = static_cast<double>(drop_index.size());
    for (int i : drop_index) {
      for (int c = 0; c < K; ++c) {
        Tree& t = *models[static_cast<size_t>(i) * K + c];
        if (!cfg.xgboost_dart_mode) {
          t.Shrinkage(1.0f / (k + 1.0f));      // (validation scores would take the tree here)
          t.Shrinkage(-k);
        } else {
          t.Shrinkage(shrinkage_rate);
          t.Shrinkage(-k / cfg.learning_rate);
        }
        AddTreeToScores(t, c);
      }
      if (!cfg.uniform_drop) {
        if (!cfg.xgboost_dart_mode) {
          sum_weight -= tree_weight[i] * (1.0f / (k + 1.0f));
          tree_weight[i] *= (k / (k + 1.0f));
        } else {
          sum_weight -= tree_weight[i] * (1.0f / (k + cfg.learning_rate));
          tree_weight[i] *= (k / (k + cfg.learning_rate));
        }
      }
    }
  }

  // ColSampler::ResetByTree
  void SampleFeatures() {
    if (!(cfg.feature_fraction < 1.0)) return;
    const int min_cnt = std::min(2, F);
    const int cnt = std::max(static_cast<int>(F * cfg.feature_fraction + 0.5), min_cnt);
    tree_mask.assign(F, 0);
    for (int f : feature_rand.Sample(F, cnt)) tree_mask[f] = 1;
  }

  // ColSampler::GetByNode (feature_fraction_bynode < 1): a fresh sample for every leaf that is searched, drawn from the tree's own
  // column sample when there is one, from all used features otherwise — on the same stream as the by-tree sample
  std::vector<char> SampleFeaturesByNode() {
    std::vector<char> mask(F, 1);
    if (!(cfg.feature_fraction_bynode < 1.0)) return mask;
    std::vector<int> pool;
    for (int f = 0; f < F; ++f) if (!(cfg.feature_fraction < 1.0) || tree_mask[f]) pool.push_back(f);
    const int total = static_cast<int>(pool.size());
    const int cnt = std::max(static_cast<int>(total * cfg.feature_fraction_bynode + 0.5), std::min(2, total));
    mask.assign(F, 0);
    for (int i : feature_rand.Sample(total, cnt)) mask[pool[i]] = 1;
    return mask;
  }

  // ---------------------------------------------------------------- objectives
  void InitObjective() {
    if (cfg.objective == "binary") {
      obj = OBJ_BINARY;
      int64_t pos = 0, neg = 0;
      for (auto& sh : shards) for (float l : sh.ds->label) { if (l > 0) ++pos; else ++neg; }   // global (summed over machines)
      need_train = !(pos == 0 || neg == 0);
      label_weights[0] = label_weights[1] = 1.0;
      if (cfg.is_unbalance && pos > 0 && neg > 0) {
        if (pos > neg) { label_weights[1] = 1.0; label_weights[0] = static_cast<double>(pos) / neg; }
        else { label_weights[1] = static_cast<double>(neg) / pos; label_weights[0] = 1.0; }
      }
      label_weights[1] *= cfg.scale_pos_weight;
    } else if (cfg.objective == "regression") {
      obj = OBJ_L2;
    } else if (cfg.objective == "huber") {
      obj = OBJ_HUBER;
    } else if (cfg.objective == "fair") {
      obj = OBJ_FAIR;
    } else if (cfg.objective == "poisson" || cfg.objective == "gamma" || cfg.objective == "tweedie") {
      obj = cfg.objective == "poisson" ? OBJ_POISSON : cfg.objective == "gamma" ? OBJ_GAMMA : OBJ_TWEEDIE;
      for (auto& sh : shards) for (float l : sh.ds->label)      // RegressionPoissonLoss::Init / Gamma: label domain
        if (l < 0.0f || (obj == OBJ_GAMMA && l <= 0.0f)) throw std::runtime_error("oracle: " + cfg.objective + " needs non-negative labels");
    } else if (cfg.objective == "regression_l1" || cfg.objective == "quantile" || cfg.objective == "mape") {
      obj = cfg.objective == "regression_l1" ? OBJ_L1 : cfg.objective == "quantile" ? OBJ_QUANTILE : OBJ_MAPE;
      if (obj == OBJ_MAPE) {      // RegressionMAPELOSS::Init
        for (auto& sh : shards) {
          const Dataset* d = sh.ds;
          std::vector<label_t> lw(d->num_data);
          for (int64_t i = 0; i < d->num_data; ++i) {
            lw[i] = 1.0f / std::max(1.0f, std::fabs(d->label[i]));
            if (has_weights) lw[i] = lw[i] * d->weight[i];
          }
          label_weight.push_back(std::move(lw));
        }
      }
    } else if (cfg.objective == "multiclass") {
      // MulticlassSoftmax::Init (upstream src/objective/multiclass_objective.hpp)
      obj = OBJ_MULTICLASS;
      K = cfg.num_class;
      if (K < 2) throw std::runtime_error("oracle: multiclass needs num_class >= 2");
      class_init_probs.assign(K, 0.0);
      double sum_weight = 0.0;
      for (auto& sh : shards) {
        const Dataset* d = sh.ds;
        for (int64_t i = 0; i < d->num_data; ++i) {
          const int l = static_cast<int>(d->label[i]);
          if (l < 0 || l >= K) throw std::runtime_error("oracle: multiclass label outside [0, num_class)");
          if (has_weights) { class_init_probs[l] += d->weight[i]; sum_weight += d->weight[i]; }
          else class_init_probs[l] += 1.0;
        }
        if (!has_weights) sum_weight += static_cast<double>(d->num_data);
      }
      class_need_train.assign(K, 1);
      for (int k = 0; k < K; ++k) {
        class_init_probs[k] /= sum_weight;
        if (std::fabs(class_init_probs[k]) <= kEpsilon || std::fabs(class_init_probs[k]) >= 1.0 - kEpsilon) class_need_train[k] = 0;
      }
      for (auto& sh : shards) {
        const size_t n = sh.score.size();
        sh.score_k.assign(K, std::vector<double>(n, 0.0));
        sh.grad_k.assign(K, std::vector<score_t>(n, 0.0f));
        sh.hess_k.assign(K, std::vector<score_t>(n, 0.0f));
        if (sh.ds->has_init_score) throw std::runtime_error("oracle: multiclass init scores are not restated");
      }
    } else if (cfg.objective == "multiclassova") {
      // MulticlassOVA::Init: one BinaryLogloss per class, positive = "label == class" (upstream src/objective/multiclass_objective.hpp)
      obj = OBJ_MULTICLASSOVA;
      K = cfg.num_class;
      if (K < 2) throw std::runtime_error("oracle: multiclassova needs num_class >= 2");
      std::vector<double> cnt(K, 0.0);
      double total = 0.0;
      ova_suml.assign(K, 0.0);
      for (auto& sh : shards) {
        const Dataset* d = sh.ds;
        for (int64_t i = 0; i < d->num_data; ++i) {
          const int l = static_cast<int>(d->label[i]);
          if (l < 0 || l >= K) throw std::runtime_error("oracle: multiclassova label outside [0, num_class)");
          const double w = has_weights ? d->weight[i] : 1.0;
          cnt[l] += 1.0; total += 1.0; ova_suml[l] += w; ova_sumw += w;
        }
      }
      class_need_train.assign(K, 1);
      ova_label_weights.assign(2 * K, 1.0);
      for (int k = 0; k < K; ++k) {
        const double pos = cnt[k], neg = total - cnt[k];
        if (pos == 0 || neg == 0) class_need_train[k] = 0;
        if (cfg.is_unbalance && pos > 0 && neg > 0) {
          if (pos > neg) ova_label_weights[2 * k] = pos / neg; else ova_label_weights[2 * k + 1] = neg / pos;
        }
        ova_label_weights[2 * k + 1] *= cfg.scale_pos_weight;
      }
      for (auto& sh : shards) {
        const size_t n = sh.score.size();
        sh.score_k.assign(K, std::vector<double>(n, 0.0));
        sh.grad_k.assign(K, std::vector<score_t>(n, 0.0f));
        sh.hess_k.assign(K, std::vector<score_t>(n, 0.0f));
        if (sh.ds->has_init_score) throw std::runtime_error("oracle: multiclass init scores are not restated");
      }
    } else if (cfg.objective == "custom" || cfg.objective == "none") {
      obj = OBJ_CUSTOM;      // no objective function: gradients come from the caller (TrainOneIter(grad, hess))
      K = std::max(1, cfg.num_class);      // num_tree_per_iteration = num_class also without an objective (GBDT::Init)
      if (K > 1) {
        class_need_train.assign(K, 1);
        for (auto& sh : shards) {
          const size_t n = sh.score.size();
          sh.score_k.assign(K, std::vector<double>(n, 0.0));
          sh.grad_k.assign(K, std::vector<score_t>(n, 0.0f));
          sh.hess_k.assign(K, std::vector<score_t>(n, 0.0f));
        }
      }
    } else if (cfg.objective == "cross_entropy") {
      obj = OBJ_XENTROPY;      // CrossEntropy (upstream src/objective/xentropy_objective.hpp): labels are probabilities in [0, 1]
      for (auto& sh : shards) for (float l : sh.ds->label) if (!(l >= 0.0f && l <= 1.0f)) throw std::runtime_error("oracle: cross_entropy needs labels in [0, 1]");
    } else if (cfg.objective == "rank_xendcg") {
      // RankXENDCG::Init (upstream src/objective/rank_objective.hpp): one random stream per query of the machine
      obj = OBJ_XENDCG;
      for (auto& sh : shards) {
        if (sh.ds->query_boundaries.size() < 2) throw std::runtime_error("oracle: rank_xendcg needs query information");
        std::vector<Random> r;
        for (size_t q = 0; q + 1 < sh.ds->query_boundaries.size(); ++q) r.emplace_back(cfg.objective_seed + static_cast<int>(q));
        xe_rands.push_back(std::move(r));
      }
    } else if (cfg.objective == "lambdarank") {
      obj = OBJ_LAMBDARANK;
      label_gain = cfg.label_gain;
      if (label_gain.empty()) { for (int i = 0; i < 31; ++i) label_gain.push_back(static_cast<double>((1LL << i) - 1)); }
      discount.resize(10000);
      for (int i = 0; i < 10000; ++i) discount[i] = 1.0 / std::log2(2.0 + i);
      inverse_max_dcgs.resize(S);
      for (int s = 0; s < S; ++s) {
        const Dataset* d = shards[s].ds;
        if (d->query_boundaries.empty()) throw std::runtime_error("lambdarank requires query information");
        int nq = static_cast<int>(d->query_boundaries.size()) - 1;
        inverse_max_dcgs[s].resize(nq);
        for (int q = 0; q < nq; ++q) {
          int b = d->query_boundaries[q], cnt = d->query_boundaries[q + 1] - b;
          double v = MaxDCGAtK(cfg.lambdarank_truncation_level, d->label.data() + b, cnt);
          if (v > 0.0) v = 1.0f / v;
          inverse_max_dcgs[s][q] = v;
        }
      }
      // sigmoid table (ConstructSigmoidTable)
      const size_t bins = 1024 * 1024;
      min_sigmoid_input = -50.0 / cfg.sigmoid / 2;
      max_sigmoid_input = -min_sigmoid_input;
      sigmoid_table.resize(bins);
      sigmoid_table_idx_factor = bins / (max_sigmoid_input - min_sigmoid_input);
      for (size_t i = 0; i < bins; ++i) {
        const double sc = i / sigmoid_table_idx_factor + min_sigmoid_input;
        sigmoid_table[i] = 1.0f / (1.0f + std::exp(sc * cfg.sigmoid));
      }
    } else {
      throw std::runtime_error("oracle: objective not restated: " + cfg.objective);
    }
  }

  double MaxDCGAtK(int k, const label_t* label, int num) const {
    double ret = 0.0;
    std::vector<int> label_cnt(label_gain.size(), 0);
    for (int i = 0; i < num; ++i) ++label_cnt[static_cast<int>(label[i])];
    int top_label = static_cast<int>(label_gain.size()) - 1;
    if (k > num) k = num;
    for (int j = 0; j < k; ++j) {
      while (top_label > 0 && label_cnt[top_label] <= 0) top_label -= 1;
      if (top_label < 0) break;
      ret += discount[j] * label_gain[top_label];
      label_cnt[top_label] -= 1;
    }
    return ret;
  }

  double GetSigmoid(double score) const {
    if (score <= min_sigmoid_input) return sigmoid_table[0];
    if (score >= max_sigmoid_input) return sigmoid_table[sigmoid_table.size() - 1];
    return sigmoid_table[static_cast<size_t>((score - min_sigmoid_input) * sigmoid_table_idx_factor)];
  }

  double BoostFromScore() const {
    double suml = 0.0, sumw = 0.0;
    if (obj == OBJ_MULTICLASS) return std::log(std::max<double>(kEpsilon, class_init_probs[cur_class]));
    if (obj == OBJ_MULTICLASSOVA) {      // BinaryLogloss::BoostFromScore of the class's one-vs-rest problem
      double pavg = ova_suml[cur_class] / ova_sumw;
      pavg = std::min(pavg, 1.0 - kEpsilon);
      pavg = std::max<double>(pavg, kEpsilon);
      return std::log(pavg / (1.0f - pavg)) / cfg.sigmoid;
    }
    if (obj == OBJ_XENTROPY) {      // CrossEntropy::BoostFromScore
      for (auto& sh : shards) {
        const Dataset* d = sh.ds;
        if (has_weights) { for (int64_t i = 0; i < d->num_data; ++i) { suml += d->label[i] * d->weight[i]; sumw += d->weight[i]; } }
        else { sumw += static_cast<double>(d->num_data); for (int64_t i = 0; i < d->num_data; ++i) suml += d->label[i]; }
      }
      double pavg = suml / sumw;
      pavg = std::min(pavg, 1.0 - kEpsilon);
      pavg = std::max<double>(pavg, kEpsilon);
      return std::log(pavg / (1.0f - pavg));
    }
    if (obj == OBJ_BINARY) {
      for (auto& sh : shards) {
        const Dataset* d = sh.ds;
        if (has_weights) { for (int64_t i = 0; i < d->num_data; ++i) { suml += (d->label[i] > 0) * d->weight[i]; sumw += d->weight[i]; } }
        else { sumw += static_cast<double>(d->num_data); for (int64_t i = 0; i < d->num_data; ++i) suml += (d->label[i] > 0); }
      }
      double pavg = suml / sumw;
      pavg = std::min(pavg, 1.0 - kEpsilon);
      pavg = std::max<double>(pavg, kEpsilon);
      return std::log(pavg / (1.0f - pavg)) / cfg.sigmoid;
    }
    if (RenewsLeaves()) {
      // RegressionL1loss / Quantile / MAPE::BoostFromScore: the (weighted) percentile of the machine's own labels; the mean over
      // machines is taken by ObtainAutomaticInitialScore (GlobalSyncUpByMean), so the caller must not average again
      double acc = 0.0;
      for (size_t s = 0; s < shards.size(); ++s) {
        const Dataset* d = shards[s].ds;
        std::vector<label_t> lab(d->label.begin(), d->label.begin() + d->num_data);
        const double a = obj == OBJ_QUANTILE ? cfg.alpha : 0.5;
        const bool af = obj == OBJ_QUANTILE;
        label_t v;
        if (obj == OBJ_MAPE) v = WeightedPercentileOf<label_t, label_t>(lab, label_weight[s], 0.5, false);
        else if (has_weights) v = WeightedPercentileOf<label_t, label_t>(lab, std::vector<label_t>(d->weight.begin(), d->weight.begin() + d->num_data), a, af);
        else v = PercentileOf<label_t>(lab, a, af);
        acc += static_cast<double>(v);
      }
      return acc / static_cast<double>(shards.size());
    }
    if (IsPointRegression()) {
      for (auto& sh : shards) {
        const Dataset* d = sh.ds;
        if (has_weights) { for (int64_t i = 0; i < d->num_data; ++i) { suml += d->label[i] * d->weight[i]; sumw += d->weight[i]; } }
        else { sumw += static_cast<double>(d->num_data); for (int64_t i = 0; i < d->num_data; ++i) suml += d->label[i]; }
      }
      const double mean = suml / sumw;
      if (LogLink()) return mean > 0.0 ? std::log(mean) : -std::numeric_limits<double>::infinity();      // Common::SafeLog
      return mean;
    }
    return 0.0;
  }

  void GetGradients() {
    if (obj == OBJ_MULTICLASS) { GetGradientsMulticlass(); return; }
    if (obj == OBJ_MULTICLASSOVA) {      // MulticlassOVA::GetGradients: BinaryLogloss::GetGradients per class on its slice of the scores
      const double sig = cfg.sigmoid;
      for (auto& sh : shards) {
        const Dataset* d = sh.ds;
        const int64_t n = d->num_data;
        for (int k = 0; k < K; ++k) {
#pragma omp parallel for schedule(static)
          for (int64_t i = 0; i < n; ++i) {
            const int is_pos = static_cast<int>(d->label[i]) == k;
            const int label = is_pos ? 1 : -1;
            const double lw = ova_label_weights[2 * k + is_pos];
            const double response = -label * sig / (1.0f + SoftExp(label * sig * sh.score_k[k][i]));
            const double abs_response = std::fabs(response);
            if (!has_weights) {
              sh.grad_k[k][i] = static_cast<score_t>(response * lw);
              sh.hess_k[k][i] = static_cast<score_t>(abs_response * (sig - abs_response) * lw);
            } else {
              sh.grad_k[k][i] = static_cast<score_t>(response * lw * d->weight[i]);
              sh.hess_k[k][i] = static_cast<score_t>(abs_response * (sig - abs_response) * lw * d->weight[i]);
            }
          }
        }
      }
      return;
    }
    for (int s = 0; s < S; ++s) {
      Shard& sh = shards[s];
      const Dataset* d = sh.ds;
      const int64_t n = d->num_data;
      if (obj == OBJ_BINARY) {
        if (!need_train) continue;
        const double sig = cfg.sigmoid;
#pragma omp parallel for schedule(static)
        for (int64_t i = 0; i < n; ++i) {
          const int is_pos = d->label[i] > 0;
          const int label = is_pos ? 1 : -1;
          const double lw = label_weights[is_pos];
          const double response = -label * sig / (1.0f + SoftExp(label * sig * sh.score[i]));
          const double abs_response = std::fabs(response);
          if (!has_weights) {
            sh.grad[i] = static_cast<score_t>(response * lw);
            sh.hess[i] = static_cast<score_t>(abs_response * (sig - abs_response) * lw);
          } else {
            sh.grad[i] = static_cast<score_t>(response * lw * d->weight[i]);
            sh.hess[i] = static_cast<score_t>(abs_response * (sig - abs_response) * lw * d->weight[i]);
          }
        }
      } else if (obj == OBJ_L2) {
#pragma omp parallel for schedule(static)
        for (int64_t i = 0; i < n; ++i) {
          if (!has_weights) {
            sh.grad[i] = static_cast<score_t>(sh.score[i] - d->label[i]);
            sh.hess[i] = 1.0f;
          } else {
            sh.grad[i] = static_cast<score_t>(static_cast<score_t>(sh.score[i] - d->label[i]) * d->weight[i]);
            sh.hess[i] = static_cast<score_t>(d->weight[i]);
          }
        }
      } else if (obj == OBJ_CUSTOM) {
        throw std::runtime_error("oracle: objective=custom needs caller-supplied gradients");
      } else if (RenewsLeaves()) {
        // RegressionL1loss / RegressionQuantileloss / RegressionMAPELOSS::GetGradients
        const score_t alpha_f = static_cast<score_t>(cfg.alpha);
        const std::vector<label_t>* lw = obj == OBJ_MAPE ? &label_weight[s] : nullptr;
#pragma omp parallel for schedule(static)
        for (int64_t i = 0; i < n; ++i) {
          if (obj == OBJ_L1) {
            const double diff = sh.score[i] - d->label[i];
            const int sg = (diff > 0) - (diff < 0);
            if (!has_weights) { sh.grad[i] = static_cast<score_t>(sg); sh.hess[i] = 1.0f; }
            else { sh.grad[i] = static_cast<score_t>(sg * d->weight[i]); sh.hess[i] = d->weight[i]; }
          } else if (obj == OBJ_QUANTILE) {
            const score_t delta = static_cast<score_t>(sh.score[i] - d->label[i]);
            const score_t g = delta >= 0 ? (1.0f - alpha_f) : -alpha_f;
            if (!has_weights) { sh.grad[i] = g; sh.hess[i] = 1.0f; }
            else { sh.grad[i] = static_cast<score_t>(g * d->weight[i]); sh.hess[i] = static_cast<score_t>(d->weight[i]); }
          } else {
            const double diff = sh.score[i] - d->label[i];
            const int sg = (diff > 0) - (diff < 0);
            sh.grad[i] = static_cast<score_t>(sg * (*lw)[i]);
            sh.hess[i] = has_weights ? d->weight[i] : 1.0f;
          }
        }
      } else if (obj == OBJ_XENTROPY) {      // CrossEntropy::GetGradients
#pragma omp parallel for schedule(static)
        for (int64_t i = 0; i < n; ++i) {
          const double z = 1.0f / (1.0f + SoftExp(-sh.score[i]));
          if (!has_weights) {
            sh.grad[i] = static_cast<score_t>(z - d->label[i]);
            sh.hess[i] = static_cast<score_t>(z * (1.0f - z));
          } else {
            sh.grad[i] = static_cast<score_t>((z - d->label[i]) * d->weight[i]);
            sh.hess[i] = static_cast<score_t>(z * (1.0f - z) * d->weight[i]);
          }
        }
      } else if (obj == OBJ_XENDCG) {
        // RankXENDCG::GetGradientsForOneQuery: cross-entropy between the softmax of the scores and a randomly perturbed ground-truth
        // distribution (2^label - uniform draw), with the first three terms of the expansion; exp is the software function both sides
        // share, the sums run in document order
        const int nq = static_cast<int>(d->query_boundaries.size()) - 1;
#pragma omp parallel for schedule(guided)
        for (int q = 0; q < nq; ++q) {
          const int start = d->query_boundaries[q];
          const int cnt = d->query_boundaries[q + 1] - start;
          score_t* lambdas = sh.grad.data() + start;
          score_t* hessians = sh.hess.data() + start;
          if (cnt <= 1) { for (int i = 0; i < cnt; ++i) { lambdas[i] = 0.0f; hessians[i] = 0.0f; } continue; }
          const double* score = sh.score.data() + start;
          std::vector<double> rho(cnt), par(cnt);
          double wmax = score[0];
          for (int i = 1; i < cnt; ++i) wmax = std::max(score[i], wmax);
          double wsum = 0.0;
          for (int i = 0; i < cnt; ++i) { rho[i] = SoftExp(score[i] - wmax); wsum += rho[i]; }
          for (int i = 0; i < cnt; ++i) rho[i] /= wsum;
          double inv_denominator = 0;
          for (int i = 0; i < cnt; ++i) {
            par[i] = static_cast<double>(1 << static_cast<int>(d->label[start + i])) - xe_rands[s][q].NextFloat();      // Phi
            inv_denominator += par[i];
          }
          inv_denominator = 1. / std::max<double>(kEpsilon, inv_denominator);
          double sum_l1 = 0.0;
          for (int i = 0; i < cnt; ++i) {
            const double term = -par[i] * inv_denominator + rho[i];
            lambdas[i] = static_cast<score_t>(term);
            par[i] = term / (1. - rho[i]);
            sum_l1 += par[i];
          }
          double sum_l2 = 0.0;
          for (int i = 0; i < cnt; ++i) {
            const double term = rho[i] * (sum_l1 - par[i]);
            lambdas[i] += static_cast<score_t>(term);
            par[i] = term / (1. - rho[i]);
            sum_l2 += par[i];
          }
          for (int i = 0; i < cnt; ++i) {
            lambdas[i] += static_cast<score_t>(rho[i] * (sum_l2 - par[i]));
            hessians[i] = static_cast<score_t>(rho[i] * (1.0 - rho[i]));
          }
          if (has_weights) {
            for (int j = 0; j < cnt; ++j) {
              lambdas[j] = static_cast<score_t>(lambdas[j] * d->weight[start + j]);
              hessians[j] = static_cast<score_t>(hessians[j] * d->weight[start + j]);
            }
          }
        }
      } else if (obj != OBJ_LAMBDARANK) {
        // regression_objective.hpp: Huber (alpha), Fair (fair_c), Poisson (poisson_max_delta_step), Gamma, Tweedie (rho); exp is the
        // software exp shared with the CUDA library
        const double alpha = cfg.alpha, c = cfg.fair_c, rho = cfg.tweedie_variance_power;
        const double exp_mds = SoftExp(cfg.poisson_max_delta_step);
#pragma omp parallel for schedule(static)
        for (int64_t i = 0; i < n; ++i) {
          const double sc = sh.score[i];
          const double y = d->label[i];
          double g, h;
          if (obj == OBJ_HUBER) {
            const double diff = sc - y;
            g = std::fabs(diff) <= alpha ? diff : ((diff > 0) - (diff < 0)) * alpha;
            h = 1.0;
          } else if (obj == OBJ_FAIR) {
            const double x = sc - y;
            g = c * x / (std::fabs(x) + c);
            h = c * c / ((std::fabs(x) + c) * (std::fabs(x) + c));
          } else if (obj == OBJ_POISSON) {
            const double e = SoftExp(sc);
            g = e - y;
            h = e * exp_mds;
          } else if (obj == OBJ_GAMMA) {
            const double e = SoftExp(-sc);
            g = 1.0 - y * e;
            h = y * e;
          } else {
            const double e1 = SoftExp((1 - rho) * sc), e2 = SoftExp((2 - rho) * sc);
            g = -y * e1 + e2;
            h = -y * (1 - rho) * e1 + (2 - rho) * e2;
          }
          if (!has_weights) { sh.grad[i] = static_cast<score_t>(g); sh.hess[i] = static_cast<score_t>(h); }
          else { sh.grad[i] = static_cast<score_t>(g * d->weight[i]); sh.hess[i] = static_cast<score_t>(h * d->weight[i]); }
        }
      } else {
        const int nq = static_cast<int>(d->query_boundaries.size()) - 1;
#pragma omp parallel for schedule(guided)
        for (int q = 0; q < nq; ++q) {
          const int start = d->query_boundaries[q];
          const int cnt = d->query_boundaries[q + 1] - start;
          LambdarankOneQuery(inverse_max_dcgs[s][q], cnt, d->label.data() + start, sh.score.data() + start,
                             sh.grad.data() + start, sh.hess.data() + start);
          if (has_weights) {
            for (int j = 0; j < cnt; ++j) {
              sh.grad[start + j] = static_cast<score_t>(sh.grad[start + j] * d->weight[start + j]);
              sh.hess[start + j] = static_cast<score_t>(sh.hess[start + j] * d->weight[start + j]);
            }
          }
        }
      }
    }
  }

  void LambdarankOneQuery(double inverse_max_dcg, int cnt, const label_t* label, const double* score, score_t* lambdas,
                          score_t* hessians) const {
    for (int i = 0; i < cnt; ++i) { lambdas[i] = 0.0f; hessians[i] = 0.0f; }
    std::vector<int> sorted_idx(cnt);
    for (int i = 0; i < cnt; ++i) sorted_idx[i] = i;
    std::stable_sort(sorted_idx.begin(), sorted_idx.end(), [score](int a, int b) { return score[a] > score[b]; });
    const double best_score = score[sorted_idx[0]];
    int worst_idx = cnt - 1;
    if (worst_idx > 0 && score[sorted_idx[worst_idx]] == kMinScore) worst_idx -= 1;
    const double worst_score = score[sorted_idx[worst_idx]];
    double sum_lambdas = 0.0;
    const double sig = cfg.sigmoid;
    for (int i = 0; i < cnt - 1 && i < cfg.lambdarank_truncation_level; ++i) {
      if (score[sorted_idx[i]] == kMinScore) continue;
      for (int j = i + 1; j < cnt; ++j) {
        if (score[sorted_idx[j]] == kMinScore) continue;
        if (label[sorted_idx[i]] == label[sorted_idx[j]]) continue;
        int high_rank, low_rank;
        if (label[sorted_idx[i]] > label[sorted_idx[j]]) { high_rank = i; low_rank = j; } else { high_rank = j; low_rank = i; }
        const int high = sorted_idx[high_rank];
        const int high_label = static_cast<int>(label[high]);
        const double high_score = score[high];
        const double high_label_gain = label_gain[high_label];
        const double high_discount = discount[high_rank];
        const int low = sorted_idx[low_rank];
        const int low_label = static_cast<int>(label[low]);
        const double low_score = score[low];
        const double low_label_gain = label_gain[low_label];
        const double low_discount = discount[low_rank];
        const double delta_score = high_score - low_score;
        const double dcg_gap = high_label_gain - low_label_gain;
        const double paired_discount = std::fabs(high_discount - low_discount);
        double delta_pair_NDCG = dcg_gap * paired_discount * inverse_max_dcg;
        if (cfg.lambdarank_norm && best_score != worst_score) delta_pair_NDCG /= (0.01f + std::fabs(delta_score));
        double p_lambda = GetSigmoid(delta_score);
        double p_hessian = p_lambda * (1.0f - p_lambda);
        p_lambda *= -sig * delta_pair_NDCG;
        p_hessian *= sig * sig * delta_pair_NDCG;
        lambdas[low] -= static_cast<score_t>(p_lambda);
        hessians[low] += static_cast<score_t>(p_hessian);
        lambdas[high] += static_cast<score_t>(p_lambda);
        hessians[high] += static_cast<score_t>(p_hessian);
        sum_lambdas -= 2 * p_lambda;
      }
    }
    if (cfg.lambdarank_norm && sum_lambdas > 0) {
      double norm_factor = SoftLog2(1 + sum_lambdas) / sum_lambdas;
      for (int i = 0; i < cnt; ++i) {
        lambdas[i] = static_cast<score_t>(lambdas[i] * norm_factor);
        hessians[i] = static_cast<score_t>(hessians[i] * norm_factor);
      }
    }
  }

  // MulticlassSoftmax::GetGradients: softmax over the row's K scores (Common::Softmax: subtract the maximum, exp, divide by
  // the sum), g = p - [label == k], h = K / (K - 1) * p * (1 - p); exp is the software function both sides share.
  void GetGradientsMulticlass() {
    const double factor = static_cast<double>(K) / (K - 1.0f);
    for (auto& sh : shards) {
      const Dataset* d = sh.ds;
      const int64_t n = d->num_data;
#pragma omp parallel for schedule(static)
      for (int64_t i = 0; i < n; ++i) {
        double rec[64];
        double wmax = sh.score_k[0][i];
        for (int k = 0; k < K; ++k) { rec[k] = sh.score_k[k][i]; wmax = std::max(rec[k], wmax); }
        double wsum = 0.0f;
        for (int k = 0; k < K; ++k) { rec[k] = SoftExp(rec[k] - wmax); wsum += rec[k]; }
        for (int k = 0; k < K; ++k) rec[k] /= static_cast<double>(wsum);
        const int label = static_cast<int>(d->label[i]);
        for (int k = 0; k < K; ++k) {
          const double p = rec[k];
          if (!has_weights) {
            sh.grad_k[k][i] = label == k ? static_cast<score_t>(p - 1.0f) : static_cast<score_t>(p);
            sh.hess_k[k][i] = static_cast<score_t>(factor * p * (1.0f - p));
          } else {
            sh.grad_k[k][i] = label == k ? static_cast<score_t>((p - 1.0f) * d->weight[i]) : static_cast<score_t>(p * d->weight[i]);
            sh.hess_k[k][i] = static_cast<score_t>(factor * p * (1.0f - p) * d->weight[i]);
          }
        }
      }
    }
  }

  // ---------------------------------------------------------------- gradient quantisation (A.6.4)
  void Discretize() {
    double max_g = 0.0, max_h = 0.0;
    bool first = true;
    for (auto& sh : shards) {
      for (size_t i = 0; i < sh.grad.size(); ++i) {
        double ag = std::fabs(sh.grad[i]), ah = std::fabs(sh.hess[i]);
        if (first) { max_g = ag; max_h = ah; first = false; }
        if (ag > max_g) max_g = ag;
        if (ah > max_h) max_h = ah;
      }
    }
    const bool const_hess = ((obj == OBJ_L2 || obj == OBJ_L1 || obj == OBJ_QUANTILE) && !has_weights) || obj == OBJ_MAPE;      // IsConstantHessian
    grad_scale = max_g / static_cast<double>(cfg.num_grad_quant_bins / 2);
    hess_scale = const_hess ? max_h : max_h / static_cast<double>(cfg.num_grad_quant_bins);
    const double inv_g = 1.0f / grad_scale, inv_h = 1.0f / hess_scale;
    for (auto& sh : shards) {
      const int64_t n = static_cast<int64_t>(sh.grad.size());
#pragma omp parallel for schedule(static)
      for (int64_t i = 0; i < n; ++i) {
        const score_t g = sh.grad[i];
        if (cfg.stochastic_rounding) {
          const uint32_t row = static_cast<uint32_t>(sh.row_offset + i);
          const float rg = HashUniform(static_cast<uint32_t>(cfg.seed), static_cast<uint32_t>(iter * K + cur_class), row, 0);
          const float rh = HashUniform(static_cast<uint32_t>(cfg.seed), static_cast<uint32_t>(iter * K + cur_class), row, 1);
          sh.qg[i] = g >= 0.0f ? static_cast<int8_t>(g * inv_g + rg) : static_cast<int8_t>(g * inv_g - rg);
          sh.qh[i] = static_cast<int8_t>(sh.hess[i] * inv_h + rh);
        } else {
          sh.qg[i] = g >= 0.0f ? static_cast<int8_t>(g * inv_g + 0.5f) : static_cast<int8_t>(g * inv_g - 0.5f);
          sh.qh[i] = static_cast<int8_t>(sh.hess[i] * inv_h + 0.5f);
        }
      }
    }
  }

  // ---------------------------------------------------------------- histograms (A.5)
  void ConstructHistogram(int leaf, const std::vector<char>& used) {
    const bool q = cfg.use_quantized_grad;
    const size_t sz = static_cast<size_t>(F) * B * 2;
    if (q) hist_i[leaf].assign(sz, 0); else hist_f[leaf].assign(sz, 0.0);
    for (int s = 0; s < S; ++s) {
      Shard& sh = shards[s];
      const Dataset* d = sh.ds;
      const data_size_t b = sh.leaf_begin[leaf], n = sh.leaf_count[leaf];
      const data_size_t* idx = sh.indices.data() + b;
      // local histogram per machine, then summed in rank order (the socket / NCCL reduce)
      std::vector<double> lf; std::vector<int64_t> li;
      if (q) li.assign(sz, 0); else lf.assign(sz, 0.0);
#pragma omp parallel for schedule(dynamic)
      for (int f = 0; f < F; ++f) {
        if (!used[f]) continue;
        const size_t col = static_cast<size_t>(f) * d->num_data;
        if (q) {
          int64_t* h = li.data() + static_cast<size_t>(f) * B * 2;
          if (d->wide) { for (data_size_t i = 0; i < n; ++i) { uint32_t bin = d->bins16[col + idx[i]]; h[2 * bin] += sh.qg[idx[i]]; h[2 * bin + 1] += sh.qh[idx[i]]; } }
          else { for (data_size_t i = 0; i < n; ++i) { uint32_t bin = d->bins8[col + idx[i]]; h[2 * bin] += sh.qg[idx[i]]; h[2 * bin + 1] += sh.qh[idx[i]]; } }
        } else {
          double* h = lf.data() + static_cast<size_t>(f) * B * 2;
          if (d->wide) { for (data_size_t i = 0; i < n; ++i) { uint32_t bin = d->bins16[col + idx[i]]; h[2 * bin] += sh.grad[idx[i]]; h[2 * bin + 1] += sh.hess[idx[i]]; } }
          else { for (data_size_t i = 0; i < n; ++i) { uint32_t bin = d->bins8[col + idx[i]]; h[2 * bin] += sh.grad[idx[i]]; h[2 * bin + 1] += sh.hess[idx[i]]; } }
        }
      }
      if (q) { for (size_t k = 0; k < sz; ++k) hist_i[leaf][k] += li[k]; }
      else if (S == 1) { hist_f[leaf].swap(lf); }
      else { for (size_t k = 0; k < sz; ++k) hist_f[leaf][k] += lf[k]; }
    }
  }

  // node_mask: the leaf's by-node column sample; a feature outside it is still searched (its splittable flag is recorded) but cannot
  // be the leaf's split (SerialTreeLearner::ComputeBestSplitForFeature: "new_split > *best_split && is_feature_used")
  void FindBestForLeaf(int leaf, const std::vector<char>& used, const std::vector<char>* node_mask = nullptr) {
    Leaf& L = leaves[leaf];
    std::vector<SplitInfo> per_feature(F);
    std::vector<char> spl(F, 0);
#pragma omp parallel for schedule(dynamic)
    for (int f = 0; f < F; ++f) {
      if (!used[f]) continue;
      const BinMapper& m = shards[0].ds->inner_mapper(f);
      SplitInfo si;
      bool ok;
      SplitParams q = sp;
      if (!mono_inner.empty()) { q.mc = true; q.cmin = L.cmin; q.cmax = L.cmax; q.mono = m.categorical ? 0 : mono_inner[f]; }
      if (m.categorical) {
        if (cfg.use_quantized_grad)
          ok = FindBestThresholdCategorical<true>(hist_i[leaf].data() + static_cast<size_t>(f) * B * 2, m, L.sum_g, L.sum_h, L.int_g, L.int_h,
                                                  grad_scale, hess_scale, L.count, q, &si);
        else
          ok = FindBestThresholdCategorical<false>(hist_f[leaf].data() + static_cast<size_t>(f) * B * 2, m, L.sum_g, L.sum_h, 0, 0, 1.0, 1.0,
                                                   L.count, q, &si);
      } else if (cfg.use_quantized_grad) {
        ok = FindBestThresholdInt(hist_i[leaf].data() + static_cast<size_t>(f) * B * 2, m, L.int_g, L.int_h, grad_scale,
                                  hess_scale, L.count, q, &si);
      } else {
        ok = FindBestThresholdFloat(hist_f[leaf].data() + static_cast<size_t>(f) * B * 2, m, L.sum_g, L.sum_h, L.count, q, &si);
      }
      si.monotone_type = q.mono;
      if (q.mono != 0) {      // LeafConstraintsBase::ComputeMonotoneSplitGainPenalty (SerialTreeLearner::ComputeBestSplitForFeature)
        const double pen = cfg.monotone_penalty, depth = L.depth;
        const double factor = pen >= depth + 1. ? kEpsilon : pen <= 1. ? 1. - pen / std::pow(2., depth) + kEpsilon : 1. - std::pow(2, pen - 1. - depth) + kEpsilon;
        si.gain *= factor;
      }
      si.feature = f;
      si.real_feature = shards[0].ds->used_features[f];
      per_feature[f] = si;
      spl[f] = ok;
    }
    SplitInfo best;
    for (int f = 0; f < F; ++f) {
      if (!used[f]) { L.splittable[f] = 0; continue; }
      L.splittable[f] = spl[f];
      if (node_mask != nullptr && !(*node_mask)[f]) continue;
      if (per_feature[f].BetterThan(best)) best = per_feature[f];
    }
    L.best = best;
  }

  // ---------------------------------------------------------------- tree learner (A.8)
  std::unique_ptr<Tree> TrainTree() {
    const int NL = cfg.num_leaves;
    const bool q = cfg.use_quantized_grad;
    if (q) Discretize();
    leaves.assign(NL, Leaf());
    for (auto& L : leaves) L.splittable.assign(F, 1);
    if (q) hist_i.resize(NL); else hist_f.resize(NL);
    SampleFeatures();
    Leaf& root = leaves[0];
    for (auto& sh : shards) {
      if (bagging) std::copy(sh.bag_indices.begin(), sh.bag_indices.begin() + sh.bag_cnt, sh.indices.begin());   // SetBaggingData
      else std::iota(sh.indices.begin(), sh.indices.end(), 0);
      std::fill(sh.leaf_begin.begin(), sh.leaf_begin.end(), 0);
      std::fill(sh.leaf_count.begin(), sh.leaf_count.end(), 0);
      sh.leaf_count[0] = sh.bag_cnt;
      root.count += sh.bag_cnt;
    }
    // root sums (over the rows in the bag)
    if (q) {
      for (auto& sh : shards) for (data_size_t i = 0; i < sh.bag_cnt; ++i) { root.int_g += sh.qg[sh.indices[i]]; root.int_h += sh.qh[sh.indices[i]]; }
      root.sum_g = static_cast<double>(root.int_g) * grad_scale;
      root.sum_h = static_cast<double>(root.int_h) * hess_scale;
    } else {
      for (auto& sh : shards) {
        double sg = 0.0, shh = 0.0;
        for (data_size_t i = 0; i < sh.bag_cnt; ++i) { sg += sh.grad[sh.indices[i]]; shh += sh.hess[sh.indices[i]]; }
        root.sum_g += sg; root.sum_h += shh;
      }
    }
    std::unique_ptr<Tree> tree(new Tree(NL));
    int left_leaf = 0, right_leaf = -1;
    for (int split = 0; split < NL - 1; ++split) {
      // BeforeFindBestSplit
      bool go = true;
      if (cfg.max_depth > 0 && tree->leaf_depth[left_leaf] >= cfg.max_depth) {
        leaves[left_leaf].best.gain = kMinScore;
        if (right_leaf >= 0) leaves[right_leaf].best.gain = kMinScore;
        go = false;
      }
      if (go) {
        data_size_t nl = leaves[left_leaf].count;
        data_size_t nr = right_leaf >= 0 ? leaves[right_leaf].count : 0;
        if (nr < cfg.min_data_in_leaf * 2 && nl < cfg.min_data_in_leaf * 2) {
          leaves[left_leaf].best.gain = kMinScore;
          if (right_leaf >= 0) leaves[right_leaf].best.gain = kMinScore;
          go = false;
        }
        if (go) {
          if (right_leaf < 0) {
            std::vector<char> used(tree_mask);
            const std::vector<char> m0 = SampleFeaturesByNode();
            ConstructHistogram(0, used);
            FindBestForLeaf(0, used, &m0);
          } else {
            int smaller = nl < nr ? left_leaf : right_leaf;
            int larger = nl < nr ? right_leaf : left_leaf;
            // the parent's histogram lives in slot left_leaf; features unsplittable at the parent stay so in both children
            std::vector<char> used(F, 1);
            for (int f = 0; f < F; ++f) used[f] = leaves[left_leaf].splittable[f] && tree_mask[f];
            if (q) { if (larger != left_leaf) hist_i[larger].swap(hist_i[left_leaf]); }
            else { if (larger != left_leaf) hist_f[larger].swap(hist_f[left_leaf]); }
            ConstructHistogram(smaller, used);
            const size_t sz = static_cast<size_t>(F) * B * 2;
            if (q) { for (size_t k = 0; k < sz; ++k) hist_i[larger][k] -= hist_i[smaller][k]; }
            else { for (size_t k = 0; k < sz; ++k) hist_f[larger][k] -= hist_f[smaller][k]; }
            const std::vector<char> ms = SampleFeaturesByNode(), ml = SampleFeaturesByNode();      // the smaller leaf draws first
            FindBestForLeaf(smaller, used, &ms);
            FindBestForLeaf(larger, used, &ml);
          }
        }
      }
      // leaf with max gain: (gain desc, feature asc, leaf asc)
      int best_leaf = 0;
      for (int l = 1; l < tree->num_leaves; ++l) if (leaves[l].best.BetterThan(leaves[best_leaf].best)) best_leaf = l;
      const SplitInfo bs = leaves[best_leaf].best;
      if (bs.gain <= 0.0) break;
      Split(tree.get(), best_leaf, bs, &left_leaf, &right_leaf);
    }
    return tree;
  }

  void Split(Tree* tree, int best_leaf, SplitInfo bs, int* left_leaf, int* right_leaf) {
    const Dataset* d0 = shards[0].ds;
    const BinMapper& m = d0->inner_mapper(bs.feature);
    *left_leaf = best_leaf;
    const int next_leaf = tree->num_leaves;
    // DataPartition::Split on every machine (stable)
    data_size_t tot_left = 0, tot_right = 0;
    for (auto& sh : shards) {
      const Dataset* d = sh.ds;
      const data_size_t b = sh.leaf_begin[best_leaf], n = sh.leaf_count[best_leaf];
      data_size_t nl = 0, nr = 0;
      const uint32_t nan_bin = static_cast<uint32_t>(m.num_bin - 1);
      for (data_size_t i = 0; i < n; ++i) {
        const data_size_t r = sh.indices[b + i];
        const uint32_t bin = d->bin_at(bs.feature, r);
        bool go_left;
        if (bs.categorical) go_left = std::find(bs.cat_threshold.begin(), bs.cat_threshold.end(), bin) != bs.cat_threshold.end();
        else if ((m.missing_type == MISSING_NAN && bin == nan_bin) || (m.missing_type == MISSING_ZERO && bin == m.default_bin)) go_left = bs.default_left;
        else go_left = bin <= bs.threshold;
        if (go_left) sh.tmp_left[nl++] = r; else sh.tmp_right[nr++] = r;
      }
      std::copy(sh.tmp_left.begin(), sh.tmp_left.begin() + nl, sh.indices.begin() + b);
      std::copy(sh.tmp_right.begin(), sh.tmp_right.begin() + nr, sh.indices.begin() + b + nl);
      sh.leaf_count[best_leaf] = nl;
      sh.leaf_begin[next_leaf] = b + nl;
      sh.leaf_count[next_leaf] = nr;
      tot_left += nl; tot_right += nr;
    }
    if (!dp) { bs.left_count = tot_left; bs.right_count = tot_right; }   // serial learner: true counts (update_cnt)
    if (bs.categorical) {      // SerialTreeLearner::SplitInner: one bitset over the bins, one over the category values
      std::vector<int> bins(bs.cat_threshold.begin(), bs.cat_threshold.end()), cats;
      for (int b : bins) cats.push_back(m.bin_2_categorical[b]);
      *right_leaf = tree->SplitCategorical(best_leaf, bs.feature, bs.real_feature, Tree::ConstructBitset(bins.data(), static_cast<int>(bins.size())),
                                           Tree::ConstructBitset(cats.data(), static_cast<int>(cats.size())), bs.left_output, bs.right_output,
                                           bs.left_count, bs.right_count, bs.left_sum_hessian, bs.right_sum_hessian,
                                           static_cast<float>(bs.gain), m.missing_type);
    } else {
      *right_leaf = tree->Split(best_leaf, bs.feature, bs.real_feature, bs.threshold, m.bin_upper_bound[bs.threshold],
                                bs.left_output, bs.right_output, bs.left_count, bs.right_count, bs.left_sum_hessian,
                                bs.right_sum_hessian, static_cast<float>(bs.gain), m.missing_type, bs.default_left);
    }
    std::vector<char> parent_spl = leaves[best_leaf].splittable;
    Leaf& L = leaves[*left_leaf];
    Leaf& R = leaves[*right_leaf];
    L.sum_g = bs.left_sum_gradient; L.sum_h = bs.left_sum_hessian; L.int_g = bs.left_int_g; L.int_h = bs.left_int_h;
    L.count = bs.left_count;
    R.sum_g = bs.right_sum_gradient; R.sum_h = bs.right_sum_hessian; R.int_g = bs.right_int_g; R.int_h = bs.right_int_h;
    R.count = bs.right_count;
    L.splittable = parent_spl; R.splittable = parent_spl;
    // BasicLeafConstraints::Update: the new leaf inherits the range; a numerical split on a constrained feature cuts both at the
    // mean of the two outputs
    R.cmin = L.cmin; R.cmax = L.cmax;
    R.depth = L.depth = L.depth + 1;
    if (!bs.categorical && bs.monotone_type != 0) {
      const double mid = (bs.left_output + bs.right_output) / 2.0f;
      if (bs.monotone_type < 0) { L.cmin = std::max(mid, L.cmin); R.cmax = std::min(mid, R.cmax); }
      else { L.cmax = std::min(mid, L.cmax); R.cmin = std::max(mid, R.cmin); }
    }
    L.best.Reset(); R.best.Reset();
  }

  // ---------------------------------------------------------------- GBDT (A.9)
  // GBDT::TrainOneIter(gradients, hessians): with caller-supplied arrays (one value per row, shard after shard) there is no
  // automatic initial score and the objective's own gradients are not computed
  bool TrainOneIter(const float* custom_grad = nullptr, const float* custom_hess = nullptr) {
    const bool custom = custom_grad != nullptr;
    const bool multi = IsMulti();
    if (obj == OBJ_CUSTOM && !custom) throw std::runtime_error("oracle: objective=custom needs caller-supplied gradients");
    if (multi && custom && obj != OBJ_CUSTOM) throw std::runtime_error("oracle: caller-supplied gradients next to a multiclass objective are not restated");
    // multiclass: class k's scores / gradients take the place of the single-class arrays while its tree is handled
    auto enter_class = [&](int k, bool with_grad) {
      cur_class = k;
      if (!multi) return;
      for (auto& sh : shards) { sh.score.swap(sh.score_k[k]); if (with_grad) { sh.grad.swap(sh.grad_k[k]); sh.hess.swap(sh.hess_k[k]); } }
    };
    auto leave_class = [&](int k, bool with_grad) {
      if (!multi) return;
      for (auto& sh : shards) { sh.score.swap(sh.score_k[k]); if (with_grad) { sh.grad.swap(sh.grad_k[k]); sh.hess.swap(sh.hess_k[k]); } }
    };
    // ObtainAutomaticInitialScore: BoostFromScore, then GlobalSyncUpByMean over machines (every machine holds the same value)
    auto automatic_init_score = [&]() {
      double v = BoostFromScore();
      if (S > 1) { double acc = 0.0; for (int s = 0; s < S; ++s) acc += v; v = acc / S; }
      return v;
    };
    if (boost == BOOST_RF) return TrainOneIterRF(enter_class, leave_class, automatic_init_score);
    // GBDT::BoostFromAverage (upstream src/boosting/gbdt.cpp), once per class
    std::vector<double> init_scores(K, 0.0);
    if (!custom) {
      for (int k = 0; k < K; ++k) {
        enter_class(k, false);
        if (models.empty() && !shards[0].ds->has_init_score && (cfg.boost_from_average || F == 0)) {
          const double v = automatic_init_score();
          if (std::fabs(v) > kEpsilon) {
            init_scores[k] = v;
            for (auto& sh : shards) for (auto& x : sh.score) x += v;
          }
        }
        leave_class(k, false);
      }
    }
    if (boost == BOOST_DART) DroppingTrees();      // DART::GetTrainingScore: before the gradients are taken
    if (custom) {
      size_t off = 0;
      for (auto& sh : shards) {      // per machine: one value per row, class after class ([num_class][num_data], as upstream)
        const size_t n = sh.grad.size();
        if (multi) {
          for (int k = 0; k < K; ++k)
            for (size_t i = 0; i < n; ++i) { sh.grad_k[k][i] = custom_grad[off + k * n + i]; sh.hess_k[k][i] = custom_hess[off + k * n + i]; }
          off += n * K;
        } else {
          for (size_t i = 0; i < n; ++i) { sh.grad[i] = custom_grad[off + i]; sh.hess[i] = custom_hess[off + i]; }
          off += n;
        }
      }
    } else {
      GetGradients();      // Boosting(): all classes at once
    }
    Bagging();
    bool should_continue = false;
    for (int k = 0; k < K; ++k) {
      enter_class(k, true);
      const bool train_this = multi ? class_need_train[k] != 0 : need_train;
      std::unique_ptr<Tree> tree(new Tree(2));
      if (train_this && F > 0) tree = TrainTree();
      if (tree->num_leaves > 1) {
        should_continue = true;
        if (RenewsLeaves()) RenewTreeOutput(tree.get());
        tree->Shrinkage(shrinkage_rate);
        for (auto& sh : shards) {
          for (int l = 0; l < tree->num_leaves; ++l) {
            const double out = tree->leaf_value[l];
            const data_size_t b = sh.leaf_begin[l], n = sh.leaf_count[l];
            for (data_size_t i = 0; i < n; ++i) sh.score[sh.indices[b + i]] += out;
          }
          // out-of-bag rows: ScoreUpdater::AddScore(tree, indices, cnt) walks the tree over the row's bins
          for (data_size_t i = sh.bag_cnt; bagging && i < static_cast<data_size_t>(sh.ds->num_data); ++i) {
            const data_size_t r = sh.bag_indices[i];
            sh.score[r] += tree->leaf_value[LeafByBins(*tree, sh.ds, r)];
          }
        }
        if (std::fabs(init_scores[k]) > kEpsilon) tree->AddBias(init_scores[k]);
      } else if (static_cast<int>(models.size()) < K) {
        // "only add default score one-time" (GBDT::TrainOneIter): with boost_from_average the initial score is already in the
        // scores (above) and only names the constant tree; without it — and with an objective and no init_score — it is obtained here
        if (!custom && obj != OBJ_CUSTOM && !cfg.boost_from_average && !shards[0].ds->has_init_score) {
          init_scores[k] = automatic_init_score();
          for (auto& sh : shards) for (auto& x : sh.score) x += init_scores[k];
        }
        tree->AsConstantTree(init_scores[k]);
      }
      models.push_back(std::move(tree));
      leave_class(k, true);
    }
    cur_class = 0;
    if (!should_continue) {
      if (static_cast<int>(models.size()) > K) for (int k = 0; k < K; ++k) models.pop_back();
      return true;
    }
    ++iter;
    if (boost == BOOST_DART) {      // DART::TrainOneIter, after GBDT::TrainOneIter (which has advanced iter_)
      Normalize();
      if (!cfg.uniform_drop) { tree_weight.push_back(shrinkage_rate); sum_weight += shrinkage_rate; }
    }
    return false;
  }

  // RF::TrainOneIter (upstream src/boosting/rf.hpp): gradients are taken ONCE, at the automatic initial scores; every tree is fit to
  // them on its own bag, carries the initial score as a bias, and the scores hold the running average of the trees.
  template <typename EnterFn, typename LeaveFn, typename InitFn>
  bool TrainOneIterRF(EnterFn enter_class, LeaveFn leave_class, InitFn automatic_init_score) {
    const bool multi = IsMulti();
    if (iter == 0 && rf_init_scores.empty()) {      // RF::Boosting
      rf_init_scores.assign(K, 0.0);
      for (int k = 0; k < K; ++k) {
        cur_class = k;
        if (!shards[0].ds->has_init_score && (cfg.boost_from_average || F == 0)) {
          const double v = automatic_init_score();
          if (std::fabs(v) > kEpsilon) rf_init_scores[k] = v;
        }
      }
      // gradients at the constant scores: the score arrays are borrowed for the call
      std::vector<std::vector<std::vector<double>>> saved_k;
      std::vector<std::vector<double>> saved;
      for (auto& sh : shards) {
        if (multi) { saved_k.push_back(sh.score_k); for (int k = 0; k < K; ++k) std::fill(sh.score_k[k].begin(), sh.score_k[k].end(), rf_init_scores[k]); }
        else { saved.push_back(sh.score); std::fill(sh.score.begin(), sh.score.end(), rf_init_scores[0]); }
      }
      GetGradients();
      for (size_t s = 0; s < shards.size(); ++s) { if (multi) shards[s].score_k = saved_k[s]; else shards[s].score = saved[s]; }
    }
    Bagging();
    for (int k = 0; k < K; ++k) {
      enter_class(k, true);
      const bool train_this = multi ? class_need_train[k] != 0 : need_train;
      std::unique_ptr<Tree> tree(new Tree(2));
      if (train_this && F > 0) tree = TrainTree();
      bool add = false;
      if (tree->num_leaves > 1) {
        if (RenewsLeaves()) { const double pred = rf_init_scores[k]; RenewTreeOutput(tree.get(), &pred); }
        if (std::fabs(rf_init_scores[k]) > kEpsilon) tree->AddBias(rf_init_scores[k]);
        add = true;
      } else if (static_cast<int>(models.size()) < K) {
        double output = 0.0;
        if (!train_this) { if (obj != OBJ_CUSTOM) output = BoostFromScore(); } else output = rf_init_scores[k];
        tree->AsConstantTree(output);
        add = true;
      }
      if (add) {      // MultiplyScore(iter), UpdateScore(tree), MultiplyScore(1 / (iter + 1))
        const double pre = static_cast<double>(iter), post = 1.0 / (iter + 1);
        for (auto& sh : shards) {
          const int64_t n = sh.ds->num_data;
#pragma omp parallel for schedule(static)
          for (int64_t i = 0; i < n; ++i) {
            double v = sh.score[i] * pre;
            v += tree->leaf_value[LeafByBins(*tree, sh.ds, static_cast<data_size_t>(i))];
            sh.score[i] = v * post;
          }
        }
      }
      models.push_back(std::move(tree));
      leave_class(k, true);
    }
    cur_class = 0;
    ++iter;
    return false;
  }

  // Tree::AddPredictionToScore's inner decision over bins (numerical): NaN bin follows default_left, else bin <= threshold
  static int LeafByBins(const Tree& t, const Dataset* d, data_size_t row) {
    if (t.num_leaves <= 1) return 0;
    int node = 0;
    while (node >= 0) {
      const int f = t.split_feature_inner[node];
      const BinMapper& m = d->inner_mapper(f);
      const uint32_t bin = d->bin_at(f, row);
      const int mt = (t.decision_type[node] >> 2) & 3;
      const bool dl = (t.decision_type[node] & 2) != 0;
      bool left;
      if (t.decision_type[node] & 1) {
        const int k = static_cast<int>(t.threshold_in_bin[node]);
        left = Tree::FindInBitset(t.cat_threshold_inner.data() + t.cat_boundaries_inner[k], t.cat_boundaries_inner[k + 1] - t.cat_boundaries_inner[k],
                                  static_cast<int>(bin));
      } else if ((mt == MISSING_NAN && bin == static_cast<uint32_t>(m.num_bin - 1)) || (mt == MISSING_ZERO && bin == m.default_bin)) left = dl;
      else left = bin <= t.threshold_in_bin[node];
      node = left ? t.left_child[node] : t.right_child[node];
    }
    return ~node;
  }

  std::string ObjectiveString() const {
    if (obj == OBJ_BINARY) return "binary sigmoid:" + FmtG(cfg.sigmoid);
    if (obj == OBJ_L2) return "regression";
    if (obj == OBJ_CUSTOM) return "";
    if (obj == OBJ_MULTICLASS) return "multiclass num_class:" + std::to_string(K);
    if (obj == OBJ_MULTICLASSOVA) return "multiclassova num_class:" + std::to_string(K) + " sigmoid:" + FmtG(cfg.sigmoid);
    if (RenewsLeaves()) return cfg.objective;
    if (obj == OBJ_XENDCG) return "rank_xendcg";
    if (obj == OBJ_XENTROPY) return "cross_entropy";
    if (obj != OBJ_LAMBDARANK) return cfg.objective;      // huber / fair / poisson / gamma / tweedie: the name alone
    return "lambdarank";
  }

  std::string SaveModelToString() const {
    const Dataset* d = shards[0].ds;
    std::string s;
    s += "tree\n";
    s += "version=v3\n";
    s += "num_class=" + std::to_string(K) + "\n";
    s += "num_tree_per_iteration=" + std::to_string(K) + "\n";
    s += "label_index=0\n";
    s += "max_feature_idx=" + std::to_string(d->num_total_features - 1) + "\n";
    if (!ObjectiveString().empty()) s += "objective=" + ObjectiveString() + "\n";      // none without an objective function
    if (boost == BOOST_RF) s += "average_output\n";
    s += "feature_names=";
    for (int j = 0; j < d->num_total_features; ++j) { if (j) s += ' '; s += d->feature_names[j]; }
    s += "\nfeature_infos=";
    for (int j = 0; j < d->num_total_features; ++j) {
      if (j) s += ' ';
      const BinMapper& m = d->mappers[j];
      if (m.is_trivial) s += "none";
      else if (m.categorical) { for (size_t b = 0; b < m.bin_2_categorical.size(); ++b) { if (b) s += ':'; s += std::to_string(m.bin_2_categorical[b]); } }   // BinMapper::bin_info_string
      else s += "[" + FmtG17(m.min_val) + ":" + FmtG17(m.max_val) + "]";
    }
    s += "\n";
    std::vector<std::string> blocks;
    for (size_t i = 0; i < models.size(); ++i) blocks.push_back("Tree=" + std::to_string(i) + "\n" + models[i]->ToString() + "\n");
    s += "tree_sizes=";
    for (size_t i = 0; i < blocks.size(); ++i) { if (i) s += ' '; s += std::to_string(blocks[i].size()); }
    s += "\n\n";
    for (auto& b : blocks) s += b;
    s += "end of trees\n\n";
    // feature importances (split counts, descending, stable by feature index)
    std::vector<int> imp(d->num_total_features, 0);
    for (auto& t : models) for (int i = 0; i < t->num_leaves - 1; ++i) imp[t->split_feature[i]]++;
    std::vector<int> order;
    for (int j = 0; j < d->num_total_features; ++j) if (imp[j] > 0) order.push_back(j);
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return imp[a] > imp[b]; });
    s += "feature_importances:\n";
    for (int j : order) s += d->feature_names[j] + "=" + std::to_string(imp[j]) + "\n";
    s += "\nparameters:\n";
    for (auto& kv : cfg.raw) s += "[" + kv.first + ": " + kv.second + "]\n";
    s += "end of parameters\n\npandas_categorical:null\n";
    return s;
  }

  void PredictRaw(const double* data, int nrow, int ncol, double* out) const {
    for (int i = 0; i < nrow; ++i) {
      double v = 0.0;
      for (auto& t : models) v += t->Predict(data + static_cast<size_t>(i) * ncol);
      out[i] = v;
    }
  }
};

}  // namespace orc
