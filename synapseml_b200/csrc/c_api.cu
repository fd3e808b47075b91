// extern "C" boundary: the LGBM_* entry points declared in include/lightgbm_b200_c_api.h.
// Nothing throws across it: every call catches, stores a thread-local message and returns -1
// (the reference turns -1 into an exception at LightGBMUtils.scala:15-20).
#include "../../include/lightgbm_b200_c_api.h"

#include <cstring>
#include <exception>
#include <memory>
#include <string>

#include "booster.h"
#include "dataset.h"
#include "network.h"

using namespace lgb;

namespace {
thread_local std::string g_last_error = "Everything is fine";
struct ByteBuffer { std::string data; };

int Fail(const char* what) {
  g_last_error = what;
  return -1;
}
}  // namespace

namespace lgb {
// lets other translation units of the library report through LGBM_GetLastError()
void SetLastErrorMessage(const char* m) { g_last_error = m; }
}  // namespace lgb

#define API_BEGIN() try {
#define API_END()                                         \
  }                                                       \
  catch (std::exception & ex) { return Fail(ex.what()); } \
  catch (...) { return Fail("unknown exception"); }      \
  return 0;

extern "C" {

const char* LGBM_GetLastError(void) { return g_last_error.c_str(); }

int LGBM_DatasetCreateFromSampledColumn(double** sample_data, int** sample_indices, int32_t ncol, const int* num_per_col,
                                        int32_t num_sample_row, int32_t num_local_row, int64_t num_dist_row,
                                        const char* parameters, DatasetHandle* out) {
  API_BEGIN();
  *out = Dataset::CreateFromSampledColumn(sample_data, sample_indices, ncol, num_per_col, num_sample_row, num_local_row,
                                          num_dist_row, parameters);
  API_END();
}

int LGBM_B200_DatasetCreateFromSampledRows(const double* rows, int32_t num_sample_row, int32_t ncol, int32_t num_local_row,
                                           int64_t num_dist_row, const char* parameters, DatasetHandle* out) {
  API_BEGIN();
  LGB_CHECK(rows != nullptr, "rows must not be null");
  *out = Dataset::CreateFromSampledColumn(nullptr, nullptr, ncol, nullptr, num_sample_row, num_local_row, num_dist_row, parameters, rows);
  API_END();
}

int LGBM_DatasetSerializeReferenceToBinary(DatasetHandle handle, ByteBufferHandle* out, int32_t* out_len) {
  API_BEGIN();
  auto bb = new ByteBuffer();
  bb->data = static_cast<Dataset*>(handle)->SerializeReference();
  *out = bb;
  *out_len = static_cast<int32_t>(bb->data.size());
  API_END();
}

int LGBM_ByteBufferGetAt(ByteBufferHandle handle, int32_t index, uint8_t* out_val) {
  API_BEGIN();
  auto bb = static_cast<ByteBuffer*>(handle);
  LGB_CHECK(index >= 0 && static_cast<size_t>(index) < bb->data.size(), "ByteBuffer index %d out of range", index);
  *out_val = static_cast<uint8_t>(bb->data[index]);
  API_END();
}

int LGBM_ByteBufferCopy(ByteBufferHandle handle, uint8_t* out, int32_t out_capacity) {
  API_BEGIN();
  auto bb = static_cast<ByteBuffer*>(handle);
  LGB_CHECK(static_cast<size_t>(out_capacity) >= bb->data.size(), "ByteBufferCopy: capacity %d < %zu", out_capacity, bb->data.size());
  memcpy(out, bb->data.data(), bb->data.size());
  API_END();
}

int LGBM_ByteBufferFree(ByteBufferHandle handle) {
  API_BEGIN();
  delete static_cast<ByteBuffer*>(handle);
  API_END();
}

int LGBM_DatasetCreateFromSerializedReference(const void* ref_buffer, int32_t ref_buffer_size, int64_t num_row,
                                              int32_t num_classes, const char* parameters, DatasetHandle* out) {
  API_BEGIN();
  (void)num_classes;
  *out = Dataset::CreateFromSerializedReference(ref_buffer, ref_buffer_size, num_row, parameters);
  API_END();
}

int LGBM_DatasetCreateByReference(const DatasetHandle reference, int64_t num_total_row, DatasetHandle* out) {
  API_BEGIN();
  *out = static_cast<const Dataset*>(reference)->CreateByReference(num_total_row);
  API_END();
}

int LGBM_DatasetInitStreaming(DatasetHandle dataset, int32_t has_weights, int32_t has_init_scores, int32_t has_queries,
                              int32_t nclasses, int32_t nthreads, int32_t omp_max_threads) {
  API_BEGIN();
  (void)omp_max_threads;
  static_cast<Dataset*>(dataset)->InitStreaming(has_weights != 0, has_init_scores != 0, has_queries != 0, nclasses, nthreads);
  API_END();
}

int LGBM_DatasetPushRowsWithMetadata(DatasetHandle dataset, const void* data, int data_type, int32_t nrow, int32_t ncol,
                                     int32_t start_row, const float* label, const float* weight, const double* init_score,
                                     const int32_t* query, int32_t tid) {
  API_BEGIN();
  static_cast<Dataset*>(dataset)->PushRows(data, data_type, nrow, ncol, start_row, label, weight, init_score, query, tid);
  API_END();
}

int LGBM_DatasetPushRowsByCSRWithMetadata(DatasetHandle dataset, const void* indptr, int indptr_type, const int32_t* indices,
                                          const void* data, int data_type, int64_t nindptr, int64_t nelem, int64_t start_row,
                                          const float* label, const float* weight, const double* init_score, const int32_t* query,
                                          int32_t tid) {
  API_BEGIN();
  static_cast<Dataset*>(dataset)->PushRowsByCSR(indptr, indptr_type, indices, data, data_type, nindptr, nelem, start_row, label, weight,
                                                init_score, query, tid);
  API_END();
}

int LGBM_DatasetSetWaitForManualFinish(DatasetHandle dataset, int wait) {
  API_BEGIN();
  static_cast<Dataset*>(dataset)->wait_for_manual_finish = wait != 0;
  API_END();
}

int LGBM_DatasetMarkFinished(DatasetHandle dataset) {
  API_BEGIN();
  static_cast<Dataset*>(dataset)->MarkFinished();
  API_END();
}

int LGBM_DatasetCreateFromMat(const void* data, int data_type, int32_t nrow, int32_t ncol, int is_row_major,
                              const char* parameters, const DatasetHandle reference, DatasetHandle* out) {
  API_BEGIN();
  LGB_CHECK(is_row_major == 1, "LGBM_DatasetCreateFromMat: only row-major input is supported (the reference always passes 1)");
  LGB_CHECK(data_type == C_API_DTYPE_FLOAT64, "LGBM_DatasetCreateFromMat: only float64 input is supported");
  std::unique_ptr<Dataset> ds;
  if (reference != nullptr) {
    ds.reset(static_cast<const Dataset*>(reference)->CreateByReference(nrow));
  } else {
    // no reference: bins are found from the first bin_construct_sample_cnt rows (the reference's samplingMode="fixed",
    // LightGBMBase.scala:744-746); laid out as SampledData does (non-zero values per column)
    Config c;
    c.Set(parameters ? parameters : "");
    const int ns = std::min<int>(nrow, c.bin_construct_sample_cnt);
    const double* d = static_cast<const double*>(data);
    std::vector<std::vector<double>> cols(ncol);
    std::vector<std::vector<int>> idxs(ncol);
    for (int i = 0; i < ns; ++i)
      for (int j = 0; j < ncol; ++j) {
        const double v = d[static_cast<size_t>(i) * ncol + j];
        if (v != 0.0) { cols[j].push_back(v); idxs[j].push_back(i); }
      }
    std::vector<double*> dp(ncol);
    std::vector<int*> ip(ncol);
    std::vector<int> cnt(ncol);
    for (int j = 0; j < ncol; ++j) { dp[j] = cols[j].data(); ip[j] = idxs[j].data(); cnt[j] = static_cast<int>(cols[j].size()); }
    ds.reset(Dataset::CreateFromSampledColumn(dp.data(), ip.data(), ncol, cnt.data(), ns, nrow, nrow, parameters));
  }
  ds->wait_for_manual_finish = true;     // labels arrive through LGBM_DatasetSetField afterwards
  ds->InitStreaming(false, false, false, 1, 1);
  ds->PushRows(data, data_type, nrow, ncol, 0, nullptr, nullptr, nullptr, nullptr, 0);
  *out = ds.release();
  API_END();
}

// Bulk mode, sparse rows (DatasetAggregator.scala: SparseAggregatedColumns.generateDataset).  Same flow as CreateFromMat: bin
// mappers from `reference` or from the first bin_construct_sample_cnt rows, then all rows through the CSR push.
int LGBM_DatasetCreateFromCSR(const void* indptr, int indptr_type, const int32_t* indices, const void* data, int data_type,
                              int64_t nindptr, int64_t nelem, int64_t num_col, const char* parameters, const DatasetHandle reference,
                              DatasetHandle* out) {
  API_BEGIN();
  LGB_CHECK(indptr_type == C_API_DTYPE_INT32 || indptr_type == C_API_DTYPE_INT64, "LGBM_DatasetCreateFromCSR: indptr must be int32 or int64");
  LGB_CHECK(data_type == C_API_DTYPE_FLOAT64 || data_type == C_API_DTYPE_FLOAT32, "LGBM_DatasetCreateFromCSR: data must be float32 or float64");
  const int64_t nrow = nindptr - 1;
  const int ncol = static_cast<int>(num_col);
  auto ptr_at = [&](int64_t i) -> int64_t {
    return indptr_type == C_API_DTYPE_INT64 ? static_cast<const int64_t*>(indptr)[i] : static_cast<int64_t>(static_cast<const int32_t*>(indptr)[i]);
  };
  auto val_at = [&](int64_t j) -> double {
    return data_type == C_API_DTYPE_FLOAT64 ? static_cast<const double*>(data)[j] : static_cast<double>(static_cast<const float*>(data)[j]);
  };
  std::unique_ptr<Dataset> ds;
  if (reference != nullptr) {
    ds.reset(static_cast<const Dataset*>(reference)->CreateByReference(nrow));
  } else {
    Config c;
    c.Set(parameters ? parameters : "");
    const int ns = static_cast<int>(std::min<int64_t>(nrow, c.bin_construct_sample_cnt));
    std::vector<std::vector<double>> cols(ncol);
    std::vector<std::vector<int>> idxs(ncol);
    for (int i = 0; i < ns; ++i)
      for (int64_t j = ptr_at(i); j < ptr_at(i + 1); ++j) {
        const int col = indices[j];
        LGB_CHECK(col >= 0 && col < ncol, "LGBM_DatasetCreateFromCSR: column index %d is outside [0, %d)", col, ncol);
        const double v = val_at(j);
        if (v != 0.0) { cols[col].push_back(v); idxs[col].push_back(i); }
      }
    std::vector<double*> dp(ncol);
    std::vector<int*> ip(ncol);
    std::vector<int> cnt(ncol);
    for (int j = 0; j < ncol; ++j) { dp[j] = cols[j].data(); ip[j] = idxs[j].data(); cnt[j] = static_cast<int>(cols[j].size()); }
    ds.reset(Dataset::CreateFromSampledColumn(dp.data(), ip.data(), ncol, cnt.data(), ns, static_cast<int>(nrow), nrow, parameters));
  }
  ds->wait_for_manual_finish = true;     // labels arrive through LGBM_DatasetSetField afterwards
  ds->InitStreaming(false, false, false, 1, 1);
  ds->PushRowsByCSR(indptr, indptr_type == C_API_DTYPE_INT64 ? 3 : 2, indices, data, data_type, nindptr, nelem, 0, nullptr, nullptr, nullptr,
                    nullptr, 0);
  *out = ds.release();
  API_END();
}

int LGBM_DatasetSetField(DatasetHandle handle, const char* field_name, const void* field_data, int num_element, int type) {
  API_BEGIN();
  auto ds = static_cast<Dataset*>(handle);
  const std::string name(field_name);
  LGB_CHECK(!ds->finished || name == "none", "LGBM_DatasetSetField(%s): the dataset is already finished", field_name);
  if (name == "label") {
    LGB_CHECK(type == C_API_DTYPE_FLOAT32 && num_element == ds->num_data, "label must be float32[num_data]");
    ds->label.assign(static_cast<const float*>(field_data), static_cast<const float*>(field_data) + num_element);
  } else if (name == "weight") {
    LGB_CHECK(type == C_API_DTYPE_FLOAT32 && num_element == ds->num_data, "weight must be float32[num_data]");
    ds->weight.assign(static_cast<const float*>(field_data), static_cast<const float*>(field_data) + num_element);
    ds->has_weight = true;
  } else if (name == "init_score") {
    LGB_CHECK(type == C_API_DTYPE_FLOAT64 && num_element >= ds->num_data && num_element % ds->num_data == 0,
              "init_score must be float64[num_data * num_class], class-major");
    ds->num_init_score_classes = static_cast<int>(num_element / ds->num_data);
    ds->init_score.assign(static_cast<const double*>(field_data), static_cast<const double*>(field_data) + num_element);
    ds->has_init_score = true;
  } else if (name == "group" || name == "query") {
    // cardinalities, DatasetUtils.scala:15-37 / LightGBMDataset.scala:171-176
    LGB_CHECK(type == C_API_DTYPE_INT32, "group must be int32 cardinalities");
    const int32_t* card = static_cast<const int32_t*>(field_data);
    ds->query_boundaries.assign(1, 0);
    for (int i = 0; i < num_element; ++i) ds->query_boundaries.push_back(ds->query_boundaries.back() + card[i]);
    LGB_CHECK(ds->query_boundaries.back() == ds->num_data, "group cardinalities sum to %d, the dataset has %lld rows",
              ds->query_boundaries.back(), (long long)ds->num_data);
    ds->has_query = true;
  } else {
    throw std::runtime_error("Unknown field name " + name);
  }
  API_END();
}

int LGBM_DatasetGetField(DatasetHandle handle, const char* field_name, int* out_len, const void** out_ptr, int* out_type) {
  API_BEGIN();
  auto ds = static_cast<Dataset*>(handle);
  const std::string name(field_name);
  if (name == "label") { *out_len = static_cast<int>(ds->label.size()); *out_ptr = ds->label.data(); *out_type = C_API_DTYPE_FLOAT32; }
  else if (name == "weight") { *out_len = static_cast<int>(ds->weight.size()); *out_ptr = ds->weight.data(); *out_type = C_API_DTYPE_FLOAT32; }
  else if (name == "init_score") { *out_len = static_cast<int>(ds->init_score.size()); *out_ptr = ds->init_score.data(); *out_type = C_API_DTYPE_FLOAT64; }
  else if (name == "group" || name == "query") { *out_len = static_cast<int>(ds->query_boundaries.size()); *out_ptr = ds->query_boundaries.data(); *out_type = C_API_DTYPE_INT32; }
  else throw std::runtime_error("Unknown field name " + name);
  API_END();
}

int LGBM_DatasetGetNumData(DatasetHandle handle, int* out) { API_BEGIN(); *out = static_cast<int>(static_cast<Dataset*>(handle)->num_data); API_END(); }
int LGBM_DatasetGetNumFeature(DatasetHandle handle, int* out) { API_BEGIN(); *out = static_cast<Dataset*>(handle)->num_total_features; API_END(); }
int LGBM_DatasetSetFeatureNames(DatasetHandle handle, const char** feature_names, int num_feature_names) {
  API_BEGIN();
  static_cast<Dataset*>(handle)->SetFeatureNames(feature_names, num_feature_names);
  API_END();
}
int LGBM_DatasetFree(DatasetHandle handle) { API_BEGIN(); delete static_cast<Dataset*>(handle); API_END(); }

int LGBM_BoosterCreate(const DatasetHandle train_data, const char* parameters, BoosterHandle* out) {
  API_BEGIN();
  auto ds = static_cast<Dataset*>(train_data);
  if (!ds->finished) ds->MarkFinished();     // bulk path: CreateFromMat + SetField, then straight to BoosterCreate
  *out = new Booster(ds, parameters);
  API_END();
}

// A failure inside an iteration of a data-parallel fit is fatal for the whole group: the other ranks are told to stop waiting
// for this one (Booster::AbortPeers) before the error goes back to the caller.
int LGBM_BoosterUpdateOneIter(BoosterHandle handle, int* is_finished) {
  API_BEGIN();
  try {
    *is_finished = static_cast<Booster*>(handle)->UpdateOneIter() ? 1 : 0;
  } catch (...) {
    static_cast<Booster*>(handle)->AbortPeers();
    throw;
  }
  API_END();
}

int LGBM_BoosterUpdateOneIterCustom(BoosterHandle handle, const float* grad, const float* hess, int* is_finished) {
  API_BEGIN();
  try {
    *is_finished = static_cast<Booster*>(handle)->UpdateOneIterCustom(grad, hess) ? 1 : 0;
  } catch (...) {
    static_cast<Booster*>(handle)->AbortPeers();
    throw;
  }
  API_END();
}

int LGBM_BoosterResetParameter(BoosterHandle handle, const char* parameters) {
  API_BEGIN();
  static_cast<Booster*>(handle)->ResetParameter(parameters);
  API_END();
}

int LGBM_BoosterSaveModelToString(BoosterHandle handle, int start_iteration, int num_iteration, int feature_importance_type,
                                  int64_t buffer_len, int64_t* out_len, char* out_str) {
  API_BEGIN();
  (void)feature_importance_type;
  const std::string s = static_cast<Booster*>(handle)->SaveModelToString(start_iteration, num_iteration);
  *out_len = static_cast<int64_t>(s.size()) + 1;
  if (*out_len <= buffer_len) memcpy(out_str, s.c_str(), *out_len);
  API_END();
}

int LGBM_BoosterDumpModel(BoosterHandle handle, int start_iteration, int num_iteration, int feature_importance_type,
                          int64_t buffer_len, int64_t* out_len, char* out_str) {
  API_BEGIN();
  const std::string s = static_cast<Booster*>(handle)->DumpModel(start_iteration, num_iteration, feature_importance_type);
  *out_len = static_cast<int64_t>(s.size()) + 1;
  if (*out_len <= buffer_len) memcpy(out_str, s.c_str(), *out_len);
  API_END();
}

int LGBM_BoosterGetNumClasses(BoosterHandle handle, int* out_len) { API_BEGIN(); *out_len = static_cast<Booster*>(handle)->NumClasses(); API_END(); }
int LGBM_BoosterGetNumFeature(BoosterHandle handle, int* out_len) { API_BEGIN(); *out_len = static_cast<Booster*>(handle)->NumTotalFeatures(); API_END(); }
int LGBM_BoosterNumModelPerIteration(BoosterHandle handle, int* out) { API_BEGIN(); *out = static_cast<Booster*>(handle)->NumClasses(); API_END(); }
int LGBM_BoosterNumberOfTotalModel(BoosterHandle handle, int* out) { API_BEGIN(); *out = static_cast<Booster*>(handle)->NumTrees(); API_END(); }
int LGBM_BoosterGetCurrentIteration(BoosterHandle handle, int* out) { API_BEGIN(); *out = static_cast<Booster*>(handle)->CurrentIteration(); API_END(); }

int LGBM_BoosterGetPredict(BoosterHandle handle, int data_idx, int64_t* out_len, double* out_result) {
  API_BEGIN();
  auto b = static_cast<Booster*>(handle);
  LGB_CHECK(!b->predict_only(), "LGBM_BoosterGetPredict: a booster loaded from a model string has no training data");
  if (data_idx > 0) {      // validation data data_idx - 1, in the order of LGBM_BoosterAddValidData
    *out_len = b->NumValidRows(data_idx - 1) * b->NumClasses();
    b->CopyValidScores(data_idx - 1, out_result);
  } else {
    LGB_CHECK(data_idx == 0, "LGBM_BoosterGetPredict: negative data_idx");
    *out_len = b->train()->num_data * b->NumClasses();      // class-major [num_class][num_data], as upstream
    b->CopyScores(out_result);
  }
  API_END();
}

int LGBM_BoosterFree(BoosterHandle handle) { API_BEGIN(); delete static_cast<Booster*>(handle); API_END(); }

int LGBM_NetworkInit(const char* machines, int local_listen_port, int listen_time_out, int num_machines) {
  API_BEGIN();
  Network::Get()->Init(machines, local_listen_port, listen_time_out, num_machines);
  API_END();
}

int LGBM_NetworkFree(void) { API_BEGIN(); Network::Get()->Free(); API_END(); }

// ------------------------------------------------------------------------------------------------ extensions
int LGBM_B200_DatasetGetFeatureInfo(DatasetHandle handle, int feature, int* num_bin, int* missing_type, int* default_bin,
                                    int* most_freq_bin, int* is_trivial, double* min_val, double* max_val) {
  API_BEGIN();
  const BinMapperH& m = static_cast<Dataset*>(handle)->mappers.at(feature);
  *num_bin = m.num_bin; *missing_type = m.missing_type; *default_bin = m.default_bin; *most_freq_bin = m.most_freq_bin;
  *is_trivial = m.is_trivial; *min_val = m.min_val; *max_val = m.max_val;
  API_END();
}
int LGBM_B200_DatasetGetBinUpperBounds(DatasetHandle handle, int feature, double* out) {
  API_BEGIN();
  const BinMapperH& m = static_cast<Dataset*>(handle)->mappers.at(feature);
  for (int i = 0; i < m.num_bin; ++i) out[i] = m.upper[i];
  API_END();
}
int LGBM_B200_DatasetGetNumUsedFeature(DatasetHandle handle, int* out) { API_BEGIN(); *out = static_cast<Dataset*>(handle)->num_logical_features(); API_END(); }
int LGBM_B200_DatasetGetNumColumns(DatasetHandle handle, int* out) { API_BEGIN(); *out = static_cast<Dataset*>(handle)->num_features(); API_END(); }
int LGBM_B200_DatasetCopyBinColumn16(DatasetHandle handle, int inner_feature, uint16_t* out) {
  API_BEGIN();
  static_cast<Dataset*>(handle)->CopyBinColumn(inner_feature, out);
  API_END();
}
int LGBM_B200_DatasetCopyBinColumn(DatasetHandle handle, int inner_feature, uint8_t* out) {
  API_BEGIN();
  auto ds = static_cast<Dataset*>(handle);
  std::vector<uint16_t> tmp(std::max<int64_t>(ds->num_data, 1));
  ds->CopyBinColumn(inner_feature, tmp.data());
  for (int64_t i = 0; i < ds->num_data; ++i) {
    LGB_CHECK(tmp[i] < 256, "feature has more than 256 bins: use LGBM_B200_DatasetCopyBinColumn16");
    out[i] = static_cast<uint8_t>(tmp[i]);
  }
  API_END();
}
int LGBM_B200_DatasetCopyRowMajor(DatasetHandle handle, int64_t row, int64_t nrows, int* out_stride, uint8_t* out) {
  API_BEGIN();
  auto ds = static_cast<Dataset*>(handle);
  *out_stride = ds->layout.rm_stride;
  if (out != nullptr) ds->CopyRowMajor(row, nrows, out);
  API_END();
}
int LGBM_B200_BoosterGetGradients(BoosterHandle handle, float* grad, float* hess) { API_BEGIN(); static_cast<Booster*>(handle)->CopyGradients(grad, hess); API_END(); }
int LGBM_B200_BoosterGetQuantGradients(BoosterHandle handle, int8_t* grad, int8_t* hess, double* scales) {
  API_BEGIN();
  static_cast<Booster*>(handle)->CopyQuantGradients(grad, hess, scales);
  API_END();
}
int LGBM_B200_BoosterRootHistogram(BoosterHandle handle, int64_t* out) {
  API_BEGIN();
  static_cast<Booster*>(handle)->DebugRootHistogram(reinterpret_cast<long long*>(out));
  API_END();
}
int LGBM_B200_BoosterTimeRootHistogram(BoosterHandle handle, int repeats, float* out_ms) {
  API_BEGIN();
  *out_ms = static_cast<Booster*>(handle)->TimeRootHistogram(repeats);
  API_END();
}
int LGBM_B200_BoosterSetHistTiming(BoosterHandle handle, int enable) { API_BEGIN(); static_cast<Booster*>(handle)->EnableHistTiming(enable != 0); API_END(); }
int LGBM_B200_DatasetTimeBinning(DatasetHandle handle, const double* rows, int32_t nrow, int repeats, float* out_ms) {
  API_BEGIN();
  *out_ms = static_cast<Dataset*>(handle)->TimeBinning(rows, nrow, repeats);
  API_END();
}
int LGBM_B200_BoosterSetProfile(BoosterHandle handle, int enable) { API_BEGIN(); static_cast<Booster*>(handle)->EnableProfile(enable != 0); API_END(); }
int LGBM_B200_BoosterGetProfile(BoosterHandle handle, int64_t buffer_len, int64_t* out_len, char* out_str) {
  API_BEGIN();
  const std::string js = static_cast<Booster*>(handle)->ProfileJson();
  *out_len = static_cast<int64_t>(js.size()) + 1;
  if (buffer_len >= *out_len) memcpy(out_str, js.c_str(), js.size() + 1);
  API_END();
}
// small host-value all-gather over the network's TCP star (CPU-only; for tests of the multi-process path)
int LGBM_B200_NetworkAllGatherInt64(int64_t mine, int64_t* out /* num_machines values */, int* out_rank, int* out_world) {
  API_BEGIN();
  std::vector<long long> all;
  Network::Get()->HostAllGatherInt64(mine, &all);
  for (size_t i = 0; i < all.size(); ++i) out[i] = all[i];
  *out_rank = Network::Get()->rank(); *out_world = Network::Get()->world();
  API_END();
}
int LGBM_B200_BoosterGetStats(BoosterHandle handle, int64_t* hist_rows_visited, double* hist_ms, int64_t* kernel_launches,
                              int64_t* hist_launches) {
  API_BEGIN();
  auto b = static_cast<Booster*>(handle);
  *hist_rows_visited = b->last_hist_rows_visited(); *hist_ms = b->last_hist_ms(); *kernel_launches = b->launches(); *hist_launches = b->hist_launches();
  API_END();
}
// Batched best-first growth, summed over all trees so far: rounds = dependent expand rounds, expansions = leaves expanded
// (partition + histogram + split search), splits = expansions that ended up in a tree, hist_kernel_launches = launches of the
// histogram kernel (a launch covers every leaf of a round; rounds enqueued past the end of a tree return at once).
int LGBM_B200_BoosterGetGrowthStats(BoosterHandle handle, int64_t* rounds, int64_t* expansions, int64_t* splits,
                                    int64_t* hist_kernel_launches) {
  API_BEGIN();
  auto b = static_cast<Booster*>(handle);
  *rounds = b->growth_rounds(); *expansions = b->growth_expansions(); *splits = b->growth_splits();
  *hist_kernel_launches = b->hist_kernel_launches();
  API_END();
}
// TCP hand-off used by LGBM_NetworkInit to distribute the NCCL unique id, exposed so the multi-process path can be tested
// without GPUs: rank 0 (first entry of `machines`) sends `blob`, every other rank receives it in place.
int LGBM_B200_NetworkBootstrapBlob(const char* machines, int local_listen_port, int listen_time_out, uint8_t* blob, int32_t bytes,
                                   int* out_rank) {
  API_BEGIN();
  auto m = Network::ParseMachines(machines ? machines : "");
  const int r = Network::FindRank(m, local_listen_port);
  Network::BroadcastBlob(m, r, listen_time_out, blob, static_cast<size_t>(bytes));
  *out_rank = r;
  API_END();
}
int LGBM_B200_BoosterTimerStart(BoosterHandle handle) { API_BEGIN(); static_cast<Booster*>(handle)->TimerStart(); API_END(); }
int LGBM_B200_BoosterTimerStop(BoosterHandle handle, float* out_ms) { API_BEGIN(); *out_ms = static_cast<Booster*>(handle)->TimerStop(); API_END(); }
int LGBM_B200_BoosterPredictRaw(BoosterHandle handle, const double* data, int32_t nrow, int32_t ncol, double* out) {
  API_BEGIN();
  int64_t n = 0;
  static_cast<Booster*>(handle)->PredictForMat(data, 1, nrow, ncol, true, 1, 0, -1, out, &n);
  API_END();
}

// ---- validation data + metrics
int LGBM_BoosterMerge(BoosterHandle handle, BoosterHandle other_handle) {
  API_BEGIN();
  LGB_CHECK(handle != nullptr && other_handle != nullptr, "LGBM_BoosterMerge: null booster handle");
  static_cast<Booster*>(handle)->MergeFrom(*static_cast<const Booster*>(other_handle));
  API_END();
}
int LGBM_BoosterAddValidData(BoosterHandle handle, const DatasetHandle valid_data) {
  API_BEGIN();
  Dataset* v = static_cast<Dataset*>(const_cast<void*>(valid_data));
  if (v != nullptr && !v->finished) v->MarkFinished();     // bulk path: CreateFromMat(reference) + SetField, as for the training data
  static_cast<Booster*>(handle)->AddValidData(v);
  API_END();
}
int LGBM_BoosterGetEvalCounts(BoosterHandle handle, int* out_len) {
  API_BEGIN();
  *out_len = static_cast<int>(static_cast<Booster*>(handle)->EvalNames().size());
  API_END();
}
int LGBM_BoosterGetEvalNames(BoosterHandle handle, const int len, int* out_len, const size_t buffer_len, size_t* out_buffer_len,
                             char** out_strs) {
  API_BEGIN();
  const std::vector<std::string> names = static_cast<Booster*>(handle)->EvalNames();
  *out_len = static_cast<int>(names.size());
  size_t need = 0;
  for (size_t i = 0; i < names.size(); ++i) {
    need = std::max(need, names[i].size() + 1);
    if (static_cast<int>(i) < len && out_strs != nullptr && buffer_len > 0) {
      const size_t c = std::min(names[i].size(), buffer_len - 1);
      memcpy(out_strs[i], names[i].data(), c);
      out_strs[i][c] = '\0';
    }
  }
  *out_buffer_len = need;
  API_END();
}
int LGBM_BoosterGetEval(BoosterHandle handle, int data_idx, int* out_len, double* out_results) {
  API_BEGIN();
  std::vector<double> v;
  static_cast<Booster*>(handle)->Eval(data_idx, &v);
  *out_len = static_cast<int>(v.size());
  for (size_t i = 0; i < v.size(); ++i) out_results[i] = v[i];
  API_END();
}

// ---- scoring
int LGBM_BoosterLoadModelFromString(const char* model_str, int* out_num_iterations, BoosterHandle* out) {
  API_BEGIN();
  if (model_str == nullptr) throw std::runtime_error("LoadModelFromString: model string is null");
  std::unique_ptr<Booster> b(new Booster(std::string(model_str)));
  if (out_num_iterations) *out_num_iterations = b->CurrentIteration();
  *out = b.release();
  API_END();
}
int LGBM_BoosterFeatureImportance(BoosterHandle handle, int num_iteration, int importance_type, double* out_results) {
  API_BEGIN();
  static_cast<Booster*>(handle)->FeatureImportance(num_iteration, importance_type, out_results);
  API_END();
}
int LGBM_BoosterCalcNumPredict(BoosterHandle handle, int num_row, int predict_type, int start_iteration, int num_iteration,
                               int64_t* out_len) {
  API_BEGIN();
  *out_len = static_cast<Booster*>(handle)->CalcNumPredict(num_row, predict_type, start_iteration, num_iteration);
  API_END();
}
namespace {
// predict_disable_shape_check=true in the prediction parameter string (LightGBMBooster.scala:526,548 always sends the key): rows with
// another number of columns than the model was trained on are accepted — columns the model does not know are ignored, columns the
// rows lack read as 0.0 (what upstream's predictor does with its zero-initialised feature buffer).
bool ShapeCheckDisabled(const char* parameter) {
  if (parameter == nullptr) return false;
  Config c;
  c.Set(parameter);
  auto it = c.raw.find("predict_disable_shape_check");
  if (it == c.raw.end()) return false;
  const std::string& v = it->second;
  return v == "true" || v == "True" || v == "1" || v == "+";
}
// rows repacked to the model's width (a host-side formatting step), row-major float64
std::vector<double> RepackRows(const void* data, int data_type, int64_t nrow, int ncol, bool row_major, int want) {
  std::vector<double> out(static_cast<size_t>(std::max<int64_t>(nrow, 0)) * want, 0.0);
  const int keep = std::min(ncol, want);
  for (int64_t i = 0; i < nrow; ++i)
    for (int j = 0; j < keep; ++j) {
      const size_t at = row_major ? static_cast<size_t>(i) * ncol + j : static_cast<size_t>(j) * nrow + i;
      out[static_cast<size_t>(i) * want + j] = data_type == C_API_DTYPE_FLOAT64 ? static_cast<const double*>(data)[at]
                                                                                 : static_cast<double>(static_cast<const float*>(data)[at]);
    }
  return out;
}
void PredictRows(Booster* b, const void* data, int data_type, int64_t nrow, int ncol, bool row_major, int predict_type, int start_iteration,
                 int num_iteration, const char* parameter, int64_t* out_len, double* out_result) {
  LGB_CHECK(data_type == C_API_DTYPE_FLOAT32 || data_type == C_API_DTYPE_FLOAT64, "Predict: data_type must be C_API_DTYPE_FLOAT32 or FLOAT64");
  const int want = b->NumTotalFeatures();
  if (ncol != want && ncol > 0 && ShapeCheckDisabled(parameter)) {
    const std::vector<double> rows = RepackRows(data, data_type, nrow, ncol, row_major, want);
    b->PredictForMat(rows.data(), C_API_DTYPE_FLOAT64, nrow, want, true, predict_type, start_iteration, num_iteration, out_result, out_len);
    return;
  }
  b->PredictForMat(data, data_type, nrow, ncol, row_major, predict_type, start_iteration, num_iteration, out_result, out_len);
}
}  // namespace

int LGBM_BoosterPredictForMat(BoosterHandle handle, const void* data, int data_type, int32_t nrow, int32_t ncol, int is_row_major,
                              int predict_type, int start_iteration, int num_iteration, const char* parameter, int64_t* out_len,
                              double* out_result) {
  API_BEGIN();
  PredictRows(static_cast<Booster*>(handle), data, data_type, nrow, ncol, is_row_major != 0, predict_type, start_iteration, num_iteration,
              parameter, out_len, out_result);
  API_END();
}
// Sparse rows for scoring (LightGBMBooster.scala:520-538, predictForCSR): the rows are expanded to dense (absent = 0.0, a host-side
// formatting step) and scored by the same GPU kernel as PredictForMat.
int LGBM_BoosterPredictForCSR(BoosterHandle handle, const void* indptr, int indptr_type, const int32_t* indices, const void* data,
                              int data_type, int64_t nindptr, int64_t nelem, int64_t num_col, int predict_type, int start_iteration,
                              int num_iteration, const char* parameter, int64_t* out_len, double* out_result) {
  API_BEGIN();
  LGB_CHECK(indptr_type == C_API_DTYPE_INT32 || indptr_type == C_API_DTYPE_INT64, "PredictForCSR: indptr must be int32 or int64");
  LGB_CHECK(data_type == C_API_DTYPE_FLOAT64 || data_type == C_API_DTYPE_FLOAT32, "PredictForCSR: data must be float32 or float64");
  auto b = static_cast<Booster*>(handle);
  const int64_t nrow = nindptr - 1;
  const int ncol = b->NumTotalFeatures();
  LGB_CHECK(num_col <= ncol, "PredictForCSR: %lld columns, the model has %d features", (long long)num_col, ncol);
  auto ptr_at = [&](int64_t i) -> int64_t {
    return indptr_type == C_API_DTYPE_INT64 ? static_cast<const int64_t*>(indptr)[i] : static_cast<int64_t>(static_cast<const int32_t*>(indptr)[i]);
  };
  LGB_CHECK(nrow <= 0 || ptr_at(nrow) - ptr_at(0) <= nelem, "PredictForCSR: indptr does not match nelem");
  std::vector<double> dense(static_cast<size_t>(std::max<int64_t>(nrow, 0)) * ncol, 0.0);
  for (int64_t i = 0; i < nrow; ++i)
    for (int64_t j = ptr_at(i); j < ptr_at(i + 1); ++j) {
      const int c = indices[j];
      LGB_CHECK(c >= 0 && c < ncol, "PredictForCSR: column index %d is outside [0, %d)", c, ncol);
      dense[static_cast<size_t>(i) * ncol + c] =
          data_type == C_API_DTYPE_FLOAT64 ? static_cast<const double*>(data)[j] : static_cast<double>(static_cast<const float*>(data)[j]);
    }
  PredictRows(b, dense.data(), C_API_DTYPE_FLOAT64, nrow, ncol, true, predict_type, start_iteration, num_iteration, parameter, out_len, out_result);
  API_END();
}
int LGBM_BoosterPredictForCSRSingleRow(BoosterHandle handle, const void* indptr, int indptr_type, const int32_t* indices,
                                       const void* data, int data_type, int64_t nindptr, int64_t nelem, int64_t num_col,
                                       int predict_type, int start_iteration, int num_iteration, const char* parameter,
                                       int64_t* out_len, double* out_result) {
  return LGBM_BoosterPredictForCSR(handle, indptr, indptr_type, indices, data, data_type, nindptr, nelem, num_col, predict_type,
                                   start_iteration, num_iteration, parameter, out_len, out_result);
}
int LGBM_BoosterPredictForMatSingleRow(BoosterHandle handle, const void* data, int data_type, int ncol, int is_row_major,
                                       int predict_type, int start_iteration, int num_iteration, const char* parameter,
                                       int64_t* out_len, double* out_result) {
  API_BEGIN();
  PredictRows(static_cast<Booster*>(handle), data, data_type, 1, ncol, is_row_major != 0, predict_type, start_iteration, num_iteration,
              parameter, out_len, out_result);
  API_END();
}

}  // extern "C"
