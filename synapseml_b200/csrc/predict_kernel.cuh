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
  const int* split_feature;   // real (column) index
  const double* threshold;
  const signed char* decision_type;
  const int* left;
  const int* right;
  const double* leaf_value;
};

constexpr int kPredictRaw = 1, kPredictNormal = 0, kPredictLeaf = 2;

// rows: element (r, c) at rows[r * row_stride + c * col_stride] (row- or column-major, float or double).
// Trees [t0, t1) of a model with K trees per iteration (tree t belongs to class t % K; t0 and t1 are multiples of K).
// out: raw / normal: [nrow][K]; leaf index: [nrow][t1 - t0].  Normal output: sigmoid_on 1 = sigmoid, 2 = exp (log-link regression),
// 3 = softmax over the row's K raw scores (Common::Softmax: subtract the maximum, exp, divide by the sum).
// average_over > 0: the model averages its iterations (boosting=rf): raw scores are divided by the number of iterations used.
template <typename T>
__global__ void k_predict(const T* __restrict__ rows, long long nrow, long long row_stride, long long col_stride, PredModelDev m, int t0,
                          int t1, int K, int average_over, int mode, int sigmoid_on, double sigmoid, double* __restrict__ out) {
  const double kZero = (double)1e-35f;
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
            double fval = static_cast<double>(row[static_cast<long long>(m.split_feature[g]) * col_stride]);
            const int dt = m.decision_type[g];
            const int mt = (dt >> 2) & 3;
            const bool is_nan = fval != fval;
            if (is_nan && mt != 2) fval = 0.0;
            bool left;
            if ((mt == 1 && fval >= -kZero && fval <= kZero) || (mt == 2 && is_nan)) left = (dt & 2) != 0;
            else left = fval <= m.threshold[g];
            node = left ? m.left[g] : m.right[g];
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

}  // namespace lgb
