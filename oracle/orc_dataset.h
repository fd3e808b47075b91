// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_common.h header; parity unpinned).
//
// Dataset construction from the driver-side sample and row binning.  Restates upstream
// DatasetLoader::ConstructFromSampleData (src/io/dataset_loader.cpp), Dataset::PushOneRow / FeatureGroup::PushData and
// Metadata query-boundary construction (src/io/metadata.cpp), SURVEY.md A.3-A.4.  Reference call sites:
//   LGBM_DatasetCreateFromSampledColumn   lightgbm/.../dataset/ReferenceDatasetUtils.scala:38-47
//   LGBM_DatasetPushRowsWithMetadata      lightgbm/.../StreamingPartitionTask.scala:220-231
//   LGBM_DatasetMarkFinished              lightgbm/.../StreamingPartitionTask.scala:140-145
//   LGBM_DatasetCreateFromMat             lightgbm/.../dataset/DatasetAggregator.scala:413-421
#pragma once
#include <memory>

#include "orc_bin.h"

namespace orc {

struct Dataset {
  Config config;
  int num_total_features = 0;
  std::vector<BinMapper> mappers;      // per real feature
  std::vector<int> used_features;      // inner index -> real feature index
  std::vector<int> real_to_inner;      // real -> inner or -1 (trivial feature)
  int64_t num_data = 0;
  int max_num_bin = 1;
  bool wide = false;                   // true: uint16 bins (some feature has > 256 bins), else uint8
  std::vector<uint8_t> bins8;          // column-major [inner_feature][num_data]
  std::vector<uint16_t> bins16;
  std::vector<float> label, weight;
  std::vector<double> init_score;
  std::vector<int32_t> query_id;       // per-row group id as pushed (streaming) ...
  std::vector<int32_t> query_boundaries;  // ... turned into boundaries when the dataset is finished
  bool has_weight = false, has_init_score = false, has_query = false;
  std::vector<std::string> feature_names;

  int num_features() const { return static_cast<int>(used_features.size()); }
  uint32_t bin_at(int inner, int64_t row) const {
    size_t o = static_cast<size_t>(inner) * num_data + row;
    return wide ? bins16[o] : bins8[o];
  }
  const BinMapper& inner_mapper(int inner) const { return mappers[used_features[inner]]; }

  // sample_data[j][k], sample_indices[j][k], num_per_col[j]: column-major non-zero sample (SampledData.scala:30-40,70-77)
  static Dataset* FromSampledColumn(double** sample_data, int** sample_indices, int ncol, const int* num_per_col,
                                    int num_sample_row, int num_local_row, int64_t num_dist_row, const char* params) {
    (void)sample_indices;
    auto ds = new Dataset();
    ds->config.Set(params ? params : "");
    const Config& c = ds->config;
    ds->num_total_features = ncol;
    ds->mappers.resize(ncol);
    // filter_cnt uses the distributed row count (num_local_row is passed as 1 by the reference, ReferenceDatasetUtils.scala:44-45)
    int64_t denom = num_dist_row > 0 ? num_dist_row : num_local_row;
    int filter_cnt = static_cast<int>(static_cast<double>(static_cast<size_t>(c.min_data_in_leaf) * static_cast<size_t>(num_sample_row)) / denom);
#pragma omp parallel for schedule(dynamic)
    for (int j = 0; j < ncol; ++j) {
      std::vector<double> vals(sample_data[j], sample_data[j] + num_per_col[j]);
      const bool cat = std::find(c.categorical_feature.begin(), c.categorical_feature.end(), j) != c.categorical_feature.end();
      const int max_bin = c.max_bin_by_feature.empty() ? c.max_bin : c.max_bin_by_feature.at(j);
      ds->mappers[j].FindBin(vals.data(), num_per_col[j], static_cast<size_t>(num_sample_row), max_bin, c.min_data_in_bin,
                             filter_cnt, c.feature_pre_filter, c.use_missing, c.zero_as_missing, cat);
    }
    ds->FinishSchema();
    ds->Allocate(num_local_row);
    return ds;
  }

  void FinishSchema() {
    used_features.clear();
    real_to_inner.assign(num_total_features, -1);
    max_num_bin = 1;
    for (int j = 0; j < num_total_features; ++j) {
      if (!mappers[j].is_trivial) {
        real_to_inner[j] = static_cast<int>(used_features.size());
        used_features.push_back(j);
        max_num_bin = std::max(max_num_bin, mappers[j].num_bin);
      }
    }
    if (feature_names.empty()) {
      for (int j = 0; j < num_total_features; ++j) feature_names.push_back("Column_" + std::to_string(j));
    }
  }

  void Allocate(int64_t n) {
    num_data = n;
    wide = max_num_bin > 256;
    if (wide) bins16.assign(static_cast<size_t>(num_features()) * n, 0);
    else bins8.assign(static_cast<size_t>(num_features()) * n, 0);
    label.assign(n, 0.0f);
  }

  Dataset* CloneSchema(int64_t n) const {
    auto ds = new Dataset();
    ds->config = config;
    ds->num_total_features = num_total_features;
    ds->mappers = mappers;
    ds->feature_names = feature_names;
    ds->FinishSchema();
    ds->Allocate(n);
    return ds;
  }

  void InitStreaming(bool has_w, bool has_init, bool has_q) {
    has_weight = has_w; has_init_score = has_init; has_query = has_q;
    if (has_w) weight.assign(num_data, 0.0f);
    if (has_init) init_score.assign(num_data, 0.0);
    if (has_q) query_id.assign(num_data, 0);
  }

  // Row-major float64 micro-batch -> column-major bins at [start_row, start_row + nrow)
  void PushRows(const double* data, int nrow, int ncol, int64_t start_row, const float* lab, const float* w,
                const double* init, const int32_t* q) {
    if (ncol != num_total_features) throw std::runtime_error("PushRows: column count mismatch");
    if (start_row < 0 || start_row + nrow > num_data) throw std::runtime_error("PushRows: row range out of bounds");
    const int nf = num_features();
#pragma omp parallel for schedule(static)
    for (int f = 0; f < nf; ++f) {
      const BinMapper& m = mappers[used_features[f]];
      const int real = used_features[f];
      const size_t off = static_cast<size_t>(f) * num_data + start_row;
      if (wide) {
        for (int i = 0; i < nrow; ++i) bins16[off + i] = static_cast<uint16_t>(m.ValueToBin(data[static_cast<size_t>(i) * ncol + real]));
      } else {
        for (int i = 0; i < nrow; ++i) bins8[off + i] = static_cast<uint8_t>(m.ValueToBin(data[static_cast<size_t>(i) * ncol + real]));
      }
    }
    if (lab) std::copy(lab, lab + nrow, label.begin() + start_row);
    if (w && has_weight) std::copy(w, w + nrow, weight.begin() + start_row);
    if (init && has_init_score) std::copy(init, init + nrow, init_score.begin() + start_row);
    if (q && has_query) std::copy(q, q + nrow, query_id.begin() + start_row);
  }

  // Runs of equal consecutive group ids become queries (the streaming path pushes one int id per row,
  // StreamingPartitionTask.scala:331-347; bulk mode sends cardinalities, DatasetUtils.scala:15-37 — same result).
  void MarkFinished() {
    if (has_query) {
      query_boundaries.clear();
      query_boundaries.push_back(0);
      for (int64_t i = 1; i < num_data; ++i) {
        if (query_id[i] != query_id[i - 1]) query_boundaries.push_back(static_cast<int32_t>(i));
      }
      query_boundaries.push_back(static_cast<int32_t>(num_data));
    }
  }

  void SetGroupCardinalities(const int32_t* card, int n) {
    has_query = true;
    query_boundaries.assign(1, 0);
    for (int i = 0; i < n; ++i) query_boundaries.push_back(query_boundaries.back() + card[i]);
    if (query_boundaries.back() != num_data) throw std::runtime_error("group cardinalities do not sum to num_data");
  }
};

}  // namespace orc
