// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_common.h header; parity unpinned).
//
// C entry points for ctypes (oracle/oracle.py).  The shapes mirror the LGBM_* functions the reference binds
//  so that tests replay one call sequence against both the oracle and the CUDA library.
#include <omp.h>

#include <cstdio>
#include <exception>

#include "orc_train.h"

using namespace orc;

static thread_local std::string g_err = "Everything is fine";
#define ORC_BEGIN try {
#define ORC_END } catch (std::exception& e) { g_err = e.what(); return -1; } catch (...) { g_err = "unknown error"; return -1; } return 0;

extern "C" {

const char* ORC_GetLastError() { return g_err.c_str(); }

double ORC_SoftExp(double x) { return SoftExp(x); }
double ORC_SoftLog2(double x) { return SoftLog2(x); }
float ORC_HashUniform(uint32_t seed, uint32_t iter, uint32_t row, uint32_t which) { return HashUniform(seed, iter, row, which); }

int ORC_FindBin(const double* values, int n, int total_sample_cnt, int max_bin, int min_data_in_bin, int min_split_data,
                int pre_filter, int use_missing, int zero_as_missing, double* out_bounds, int* out_num_bin,
                int* out_missing_type, int* out_default_bin, int* out_most_freq_bin, int* out_trivial) {
  ORC_BEGIN
  std::vector<double> v(values, values + n);
  BinMapper m;
  m.FindBin(v.data(), n, static_cast<size_t>(total_sample_cnt), max_bin, min_data_in_bin, min_split_data, pre_filter != 0,
            use_missing != 0, zero_as_missing != 0);
  for (int i = 0; i < m.num_bin; ++i) out_bounds[i] = m.bin_upper_bound[i];
  *out_num_bin = m.num_bin; *out_missing_type = m.missing_type; *out_default_bin = m.default_bin;
  *out_most_freq_bin = m.most_freq_bin; *out_trivial = m.is_trivial;
  ORC_END
}

int ORC_DatasetCreateFromSampledColumn(double** sample_data, int** sample_indices, int32_t ncol, const int* num_per_col,
                                       int32_t num_sample_row, int32_t num_local_row, int64_t num_dist_row,
                                       const char* params, void** out) {
  ORC_BEGIN
  *out = Dataset::FromSampledColumn(sample_data, sample_indices, ncol, num_per_col, num_sample_row, num_local_row,
                                    num_dist_row, params);
  ORC_END
}

int ORC_DatasetCreateByReference(void* ref, int64_t num_rows, void** out) {
  ORC_BEGIN
  *out = static_cast<Dataset*>(ref)->CloneSchema(num_rows);
  ORC_END
}

int ORC_DatasetInitStreaming(void* h, int32_t has_w, int32_t has_init, int32_t has_q, int32_t, int32_t, int32_t) {
  ORC_BEGIN
  static_cast<Dataset*>(h)->InitStreaming(has_w != 0, has_init != 0, has_q != 0);
  ORC_END
}

int ORC_DatasetPushRowsWithMetadata(void* h, const void* data, int data_type, int32_t nrow, int32_t ncol, int32_t start_row,
                                    const float* label, const float* weight, const double* init_score, const int32_t* query,
                                    int32_t) {
  ORC_BEGIN
  if (data_type != 1) throw std::runtime_error("oracle: only C_API_DTYPE_FLOAT64 is restated");
  static_cast<Dataset*>(h)->PushRows(static_cast<const double*>(data), nrow, ncol, start_row, label, weight, init_score, query);
  ORC_END
}

int ORC_DatasetMarkFinished(void* h) { ORC_BEGIN static_cast<Dataset*>(h)->MarkFinished(); ORC_END }

int ORC_DatasetSetGroup(void* h, const int32_t* card, int n) { ORC_BEGIN static_cast<Dataset*>(h)->SetGroupCardinalities(card, n); ORC_END }

int ORC_DatasetGetNumData(void* h, int* out) { ORC_BEGIN *out = static_cast<int>(static_cast<Dataset*>(h)->num_data); ORC_END }
int ORC_DatasetGetNumFeature(void* h, int* out) { ORC_BEGIN *out = static_cast<Dataset*>(h)->num_total_features; ORC_END }
int ORC_DatasetGetNumUsedFeature(void* h, int* out) { ORC_BEGIN *out = static_cast<Dataset*>(h)->num_features(); ORC_END }

int ORC_DatasetSetFeatureNames(void* h, const char** names, int n) {
  ORC_BEGIN
  auto d = static_cast<Dataset*>(h);
  d->feature_names.assign(names, names + n);
  ORC_END
}

int ORC_DatasetGetFeatureInfo(void* h, int real_feature, int* num_bin, int* missing_type, int* default_bin,
                              int* most_freq_bin, int* is_trivial, double* min_val, double* max_val) {
  ORC_BEGIN
  const BinMapper& m = static_cast<Dataset*>(h)->mappers.at(real_feature);
  *num_bin = m.num_bin; *missing_type = m.missing_type; *default_bin = m.default_bin; *most_freq_bin = m.most_freq_bin;
  *is_trivial = m.is_trivial; *min_val = m.min_val; *max_val = m.max_val;
  ORC_END
}

int ORC_DatasetGetBinUpperBounds(void* h, int real_feature, double* out) {
  ORC_BEGIN
  const BinMapper& m = static_cast<Dataset*>(h)->mappers.at(real_feature);
  for (int i = 0; i < m.num_bin; ++i) out[i] = m.categorical ? static_cast<double>(m.bin_2_categorical[i]) : m.bin_upper_bound[i];   // categorical: the category of every bin
  ORC_END
}

// bins of one used (inner) feature for rows [0, num_data) as uint16
int ORC_DatasetGetBinColumn(void* h, int inner_feature, uint16_t* out) {
  ORC_BEGIN
  auto d = static_cast<Dataset*>(h);
  for (int64_t i = 0; i < d->num_data; ++i) out[i] = static_cast<uint16_t>(d->bin_at(inner_feature, i));
  ORC_END
}

int ORC_DatasetFree(void* h) { ORC_BEGIN delete static_cast<Dataset*>(h); ORC_END }

int ORC_BoosterCreate(void** datasets, int n_shards, const char* params, void** out) {
  ORC_BEGIN
  std::vector<const Dataset*> ds;
  for (int i = 0; i < n_shards; ++i) ds.push_back(static_cast<const Dataset*>(datasets[i]));
  *out = new Booster(ds, params);
  ORC_END
}

int ORC_BoosterUpdateOneIter(void* h, int* is_finished) {
  ORC_BEGIN
  *is_finished = static_cast<Booster*>(h)->TrainOneIter() ? 1 : 0;
  ORC_END
}

int ORC_BoosterUpdateOneIterCustom(void* h, const float* grad, const float* hess, int* is_finished) {
  ORC_BEGIN
  *is_finished = static_cast<Booster*>(h)->TrainOneIter(grad, hess) ? 1 : 0;
  ORC_END
}

int ORC_BoosterSaveModelToString(void* h, int64_t buf_len, int64_t* out_len, char* out) {
  ORC_BEGIN
  std::string s = static_cast<Booster*>(h)->SaveModelToString();
  *out_len = static_cast<int64_t>(s.size()) + 1;
  if (*out_len <= buf_len) std::memcpy(out, s.c_str(), *out_len);
  ORC_END
}

int ORC_BoosterGetScores(void* h, int shard, double* out) {
  ORC_BEGIN
  auto& sh = static_cast<Booster*>(h)->shards.at(shard);
  if (sh.score_k.empty()) std::copy(sh.score.begin(), sh.score.end(), out);
  else for (size_t k = 0; k < sh.score_k.size(); ++k) std::copy(sh.score_k[k].begin(), sh.score_k[k].end(), out + k * sh.score_k[k].size());   // class-major
  ORC_END
}

int ORC_BoosterGetGradients(void* h, int shard, float* g, float* hh) {
  ORC_BEGIN
  auto& sh = static_cast<Booster*>(h)->shards.at(shard);
  if (sh.grad_k.empty()) {
    std::copy(sh.grad.begin(), sh.grad.end(), g);
    std::copy(sh.hess.begin(), sh.hess.end(), hh);
  } else {
    for (size_t k = 0; k < sh.grad_k.size(); ++k) {      // class-major
      std::copy(sh.grad_k[k].begin(), sh.grad_k[k].end(), g + k * sh.grad_k[k].size());
      std::copy(sh.hess_k[k].begin(), sh.hess_k[k].end(), hh + k * sh.hess_k[k].size());
    }
  }
  ORC_END
}

int ORC_BoosterGetQuantGradients(void* h, int shard, int8_t* g, int8_t* hh, double* scales) {
  ORC_BEGIN
  auto b = static_cast<Booster*>(h);
  auto& sh = b->shards.at(shard);
  std::copy(sh.qg.begin(), sh.qg.end(), g);
  std::copy(sh.qh.begin(), sh.qh.end(), hh);
  scales[0] = b->grad_scale; scales[1] = b->hess_scale;
  ORC_END
}

int ORC_BoosterNumTrees(void* h, int* out) { ORC_BEGIN *out = static_cast<int>(static_cast<Booster*>(h)->models.size()); ORC_END }

int ORC_BoosterPredictRaw(void* h, const double* data, int nrow, int ncol, double* out) {
  ORC_BEGIN
  static_cast<Booster*>(h)->PredictRaw(data, nrow, ncol, out);
  ORC_END
}

int ORC_BoosterFree(void* h) { ORC_BEGIN delete static_cast<Booster*>(h); ORC_END }

int ORC_SetNumThreads(int n) { ORC_BEGIN if (n > 0) omp_set_num_threads(n); ORC_END }
int ORC_GetMaxThreads() { return omp_get_max_threads(); }

}  // extern "C"

extern "C" int ORC_BoosterNumClass(void* h, int* out) { ORC_BEGIN *out = static_cast<orc::Booster*>(h)->K; ORC_END }
