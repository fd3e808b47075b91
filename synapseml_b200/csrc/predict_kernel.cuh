// Tree-ensemble prediction on raw feature values (the scoring side of the model text: SURVEY.md §8(f) rank 1).
// Replaces lightgbmlib's LGBM_BoosterPredictForMat / ...ForMatSingle as called from
// lightgbm/src/main/scala/com/microsoft/azure/synapse/ml/lightgbm/booster/LightGBMBooster.scala:520-557.
//
// One thread per row walks the trees in model order and adds the leaf values in that order, so the raw score of a row is
// the same double whatever the batch size (and equal to a sequential host traversal).  The decision at a node is upstream's
// NumericalDecision: NaN counts as 0.0 unless the node's missing type is NaN; a missing value (zero for missing type Zero,
// NaN for missing type NaN) follows default_left; everything else compares value <= threshold.
#pragma once
#include <cstdint>

#include "softmath.cuh"

namespace lgb {

struct PredModelDev {
  const int* node_off;        // [ntrees + 1] first internal node of every tree in the flat arrays
  const int* leaf_off;        // [ntrees + 1]
  const int* internal_count;  // rows of the training data below every internal node (feature contributions only)
  const int* leaf_count;
  const double* expected;     // [ntrees] Tree::ExpectedValue(): the count-weighted mean leaf value
  const int* max_depth;       // [ntrees]
  const int* split_feature;   // real (column) index
  const double* threshold;
  const signed char* decision_type;
  const int* left;
  const int* right;
  const double* leaf_value;
  const int* cat_begin;              // categorical nodes (decision_type bit 0): the node's category set is words
  const int* cat_words;              //   [cat_begin, cat_begin + cat_words) of cat_threshold, one bit per category value
  const unsigned int* cat_threshold;
};

// Tree::Decision on a raw value.  Numerical: NaN counts as 0.0 unless the node's missing type is NaN; a missing value (zero for
// missing type Zero, NaN for missing type NaN) follows default_left; everything else compares value <= threshold.  Categorical
// (Tree::CategoricalDecision): the value is cut to an int; NaN, negative and unlisted values go right.
__device__ __forceinline__ bool PredGoesLeft(const PredModelDev& m, int g, double fval) {
  const double kZero = (double)1e-35f;
  const int dt = m.decision_type[g];
  if (dt & 1) {
    if (fval != fval || !(fval < 2147483648.0)) return false;
    const int iv = static_cast<int>(fval);
    if (iv < 0) return false;
    const int w = iv >> 5;
    if (w >= m.cat_words[g]) return false;
    return ((m.cat_threshold[m.cat_begin[g] + w] >> (iv & 31)) & 1u) != 0u;
  }
  const int mt = (dt >> 2) & 3;
  const bool is_nan = fval != fval;
  if (is_nan && mt != 2) fval = 0.0;
  if ((mt == 1 && fval >= -kZero && fval <= kZero) || (mt == 2 && is_nan)) return (dt & 2) != 0;
  return fval <= m.threshold[g];
}

constexpr int kPredictRaw = 1, kPredictNormal = 0, kPredictLeaf = 2, kPredictContrib = 3;

// rows: element (r, c) at rows[r * row_stride + c * col_stride] (row- or column-major, float or double).
// Trees [t0, t1) of a model with K trees per iteration (tree t belongs to class t % K; t0 and t1 are multiples of K).
// out: raw / normal: [nrow][K]; leaf index: [nrow][t1 - t0].  Normal output: sigmoid_on 1 = sigmoid, 2 = exp (log-link regression),
// 3 = softmax over the row's K raw scores (Common::Softmax: subtract the maximum, exp, divide by the sum).
// average_over > 0: the model averages its iterations (boosting=rf): raw scores are divided by the number of iterations used.
template <typename T>
__global__ void k_predict(const T* __restrict__ rows, long long nrow, long long row_stride, long long col_stride, PredModelDev m, int t0,
                          int t1, int K, int average_over, int mode, int sigmoid_on, double sigmoid, double* __restrict__ out) {
  for (long long r = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; r < nrow; r += static_cast<long long>(gridDim.x) * blockDim.x) {
    const T* row = rows + r * row_stride;
    for (int cls = 0; cls < K; ++cls) {
      double acc = 0.0;
      for (int t = t0 + cls; t < t1; t += K) {
        const int no = m.node_off[t], lo = m.leaf_off[t];
        int leaf = 0;
        if (m.node_off[t + 1] > no) {
          int node = 0;
          do {
            const int g = no + node;
            const double fval = static_cast<double>(row[static_cast<long long>(m.split_feature[g]) * col_stride]);
            node = PredGoesLeft(m, g, fval) ? m.left[g] : m.right[g];
          } while (node >= 0);
          leaf = ~node;
        }
        if (mode == kPredictLeaf) out[r * (t1 - t0) + (t - t0)] = static_cast<double>(leaf);
        else acc += m.leaf_value[lo + leaf];
      }
      if (mode != kPredictLeaf) {
        if (average_over > 0) acc /= average_over;      // average_output (random forest): the mean over the iterations used
        if (mode == kPredictNormal && sigmoid_on == 1) acc = 1.0 / (1.0 + SoftExp(-sigmoid * acc));
        else if (mode == kPredictNormal && sigmoid_on == 2) acc = SoftExp(acc);      // log-link regression objectives
        out[r * K + cls] = acc;
      }
    }
    if (mode == kPredictNormal && sigmoid_on == 3) {
      double* o = out + r * K;
      double wmax = o[0];
      for (int k = 1; k < K; ++k) wmax = fmax(wmax, o[k]);
      double wsum = 0.0;
      for (int k = 0; k < K; ++k) { o[k] = SoftExp(o[k] - wmax); wsum += o[k]; }
      for (int k = 0; k < K; ++k) o[k] /= wsum;
    }
  }
}

// ------------------------------------------------------------------------------------------------ feature contributions
// predict_type 3 (C_API_PREDICT_CONTRIB; LightGBMBooster.scala:148,418-427): TreeSHAP (Lundberg et al.; upstream Tree::TreeSHAP /
// ExtendPath / UnwindPath / UnwoundPathSum in src/io/tree.cpp), the exact Shapley values of a tree's output under the tree's own
// cover counts.  One thread per row walks the trees in model order and adds into its row of the output, [K][ncol + 1] per row with
// the expected value in the last column, so a row's contributions sum to its raw score.  The recursion keeps the unique feature
// path of the current node in the thread's slice of `paths` ((D + 1)(D + 2) / 2 elements for maximum depth D).
struct ShapPathElement { int feature_index; double zero_fraction, one_fraction, pweight; };

__device__ __forceinline__ void ShapExtendPath(ShapPathElement* up, int depth, double zero_fraction, double one_fraction, int feature_index) {
  up[depth].feature_index = feature_index;
  up[depth].zero_fraction = zero_fraction;
  up[depth].one_fraction = one_fraction;
  up[depth].pweight = depth == 0 ? 1.0 : 0.0;
  for (int i = depth - 1; i >= 0; --i) {
    up[i + 1].pweight += one_fraction * up[i].pweight * (i + 1) / static_cast<double>(depth + 1);
    up[i].pweight = zero_fraction * up[i].pweight * (depth - i) / static_cast<double>(depth + 1);
  }
}
__device__ __forceinline__ void ShapUnwindPath(ShapPathElement* up, int depth, int path_index) {
  const double one_fraction = up[path_index].one_fraction, zero_fraction = up[path_index].zero_fraction;
  double next_one_portion = up[depth].pweight;
  for (int i = depth - 1; i >= 0; --i) {
    if (one_fraction != 0) {
      const double tmp = up[i].pweight;
      up[i].pweight = next_one_portion * (depth + 1) / static_cast<double>((i + 1) * one_fraction);
      next_one_portion = tmp - up[i].pweight * zero_fraction * (depth - i) / static_cast<double>(depth + 1);
    } else {
      up[i].pweight = (up[i].pweight * (depth + 1)) / static_cast<double>(zero_fraction * (depth - i));
    }
  }
  for (int i = path_index; i < depth; ++i) {
    up[i].feature_index = up[i + 1].feature_index;
    up[i].zero_fraction = up[i + 1].zero_fraction;
    up[i].one_fraction = up[i + 1].one_fraction;
  }
}
__device__ __forceinline__ double ShapUnwoundPathSum(const ShapPathElement* up, int depth, int path_index) {
  const double one_fraction = up[path_index].one_fraction, zero_fraction = up[path_index].zero_fraction;
  double next_one_portion = up[depth].pweight;
  double total = 0;
  for (int i = depth - 1; i >= 0; --i) {
    if (one_fraction != 0) {
      const double tmp = next_one_portion * (depth + 1) / static_cast<double>((i + 1) * one_fraction);
      total += tmp;
      next_one_portion = up[i].pweight - tmp * zero_fraction * ((depth - i) / static_cast<double>(depth + 1));
    } else {
      total += (up[i].pweight / zero_fraction) / ((depth - i) / static_cast<double>(depth + 1));
    }
  }
  return total;
}

template <typename T>
struct ShapCtx {
  const T* row; long long col_stride;
  const PredModelDev* m; int no, lo;      // the tree's first internal node / first leaf in the flat arrays
  double* phi;
};

template <typename T>
__device__ void TreeShapRecurse(const ShapCtx<T>& c, int node, int depth, ShapPathElement* parent_path, double parent_zero_fraction,
                                double parent_one_fraction, int parent_feature_index) {
  ShapPathElement* up = parent_path + depth;
  for (int i = 0; i < depth; ++i) up[i] = parent_path[i];
  ShapExtendPath(up, depth, parent_zero_fraction, parent_one_fraction, parent_feature_index);
  const PredModelDev& m = *c.m;
  if (node < 0) {
    const double leaf = m.leaf_value[c.lo + ~node];
    for (int i = 1; i <= depth; ++i) {
      const double w = ShapUnwoundPathSum(up, depth, i);
      c.phi[up[i].feature_index] += w * (up[i].one_fraction - up[i].zero_fraction) * leaf;
    }
    return;
  }
  const int g = c.no + node;
  const int feat = m.split_feature[g];
  const bool left = PredGoesLeft(m, g, static_cast<double>(c.row[static_cast<long long>(feat) * c.col_stride]));
  const int hot = left ? m.left[g] : m.right[g];
  const int cold = left ? m.right[g] : m.left[g];
  auto count = [&](int nd) { return static_cast<double>(nd >= 0 ? m.internal_count[c.no + nd] : m.leaf_count[c.lo + ~nd]); };
  const double w = count(node);
  const double hot_zero_fraction = count(hot) / w, cold_zero_fraction = count(cold) / w;
  double incoming_zero_fraction = 1, incoming_one_fraction = 1;
  // a feature the path has already split on is taken out and re-entered with this node's fractions
  int path_index = 0;
  for (; path_index <= depth; ++path_index)
    if (up[path_index].feature_index == feat) break;
  if (path_index != depth + 1) {
    incoming_zero_fraction = up[path_index].zero_fraction;
    incoming_one_fraction = up[path_index].one_fraction;
    ShapUnwindPath(up, depth, path_index);
    depth -= 1;
  }
  TreeShapRecurse(c, hot, depth + 1, up, hot_zero_fraction * incoming_zero_fraction, incoming_one_fraction, feat);
  TreeShapRecurse(c, cold, depth + 1, up, cold_zero_fraction * incoming_zero_fraction, 0.0, feat);
}

// out: [nrow][K][ncol + 1], zero on entry.  paths: [threads of the grid][path_elems].
template <typename T>
__global__ void k_predict_contrib(const T* __restrict__ rows, long long nrow, long long row_stride, long long col_stride, PredModelDev m,
                                  int t0, int t1, int K, int ncol, int average_over, ShapPathElement* __restrict__ paths, int path_elems,
                                  double* __restrict__ out) {
  const long long tid = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  ShapPathElement* my_paths = paths + tid * path_elems;
  for (long long r = tid; r < nrow; r += static_cast<long long>(gridDim.x) * blockDim.x) {
    for (int t = t0; t < t1; ++t) {
      double* phi = out + (r * K + (t - t0) % K) * (ncol + 1);
      phi[ncol] += m.expected[t];
      if (m.node_off[t + 1] > m.node_off[t]) {
        ShapCtx<T> c{rows + r * row_stride, col_stride, &m, m.node_off[t], m.leaf_off[t], phi};
        TreeShapRecurse(c, 0, 0, my_paths, 1.0, 1.0, -1);
      }
    }
    if (average_over > 0)
      for (int i = 0; i < K * (ncol + 1); ++i) out[r * K * (ncol + 1) + i] /= average_over;
  }
}

}  // namespace lgb
