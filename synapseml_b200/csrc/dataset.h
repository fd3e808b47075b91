// Dataset: bin mappers found on the GPU from the driver-side sample (K1), the streaming row-binning path (K2) and the
// per-row metadata.  Replaces what lightgbmlib does behind these reference call sites:
//   LGBM_DatasetCreateFromSampledColumn        lightgbm/.../dataset/ReferenceDatasetUtils.scala:38-47
//   LGBM_DatasetSerializeReferenceToBinary     lightgbm/.../dataset/ReferenceDatasetUtils.scala:52-55
//   LGBM_DatasetCreateFromSerializedReference  lightgbm/.../dataset/ReferenceDatasetUtils.scala:121-138
//   LGBM_DatasetInitStreaming                  lightgbm/.../dataset/ReferenceDatasetUtils.scala:87-94
//   LGBM_DatasetPushRowsWithMetadata           lightgbm/.../StreamingPartitionTask.scala:220-231
//   LGBM_DatasetMarkFinished                   lightgbm/.../StreamingPartitionTask.scala:140-145
//   LGBM_DatasetCreateFromMat                  lightgbm/.../dataset/DatasetAggregator.scala:413-421
#pragma once
#include <deque>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "config.h"
#include "util.h"

namespace lgb {

enum MissingType { kMissingNone = 0, kMissingZero = 1, kMissingNaN = 2 };

constexpr int kMaxBinsPerFeature = 256;   // uint8 bin matrix (uint16 storage is not built yet; rejected loudly)

struct BinMapperH {
  int num_bin = 1;
  int missing_type = kMissingNone;
  int is_trivial = 1;
  int default_bin = 0;
  int most_freq_bin = 0;
  double sparse_rate = 1.0;
  double min_val = 0.0, max_val = 0.0;
  std::vector<double> upper;    // bin_upper_bound[num_bin]; categorical: the category of every bin as a double
  int bin_type = 0;             // 0 numerical, 1 categorical (BinType::CategoricalBin)
  std::vector<int> bin_2_cat;   // categorical: category of bin b (bin 0: -1 = NaN, negative, rare and unseen categories)
};

// Per used (inner) feature, device side.
struct FeatureMetaDev {
  int num_bin;
  int missing_type;
  int default_bin;
  int most_freq_bin;
  int real_index;
  int bin_type;        // 1: categorical
  int pad_[2];
};

// HBM layout of the binned matrix.  Two copies are kept, both written by the K2 kernel in one pass:
//   cm: column-major  [F_used][n_pad]  uint8 — one byte per (feature,row); read by the row-partition kernel (one column).
//   rm: row-major     [N][rm_stride]   uint8 — one record per row; read by the histogram kernel (gathers by row index
//       fetch whole 32-byte sectors, so a row's features must be contiguous).
struct BinLayout {
  int rows_per_step;     // R: rows one warp-wide load covers in the histogram kernel
  int tile_features_q;   // features per histogram tile in quantized mode  (32/R * 4)
  int tile_features_f;   // features per histogram tile in float mode      (32/R * 2)
  int rm_stride;         // bytes per row record
};

class Dataset {
 public:
  Config cfg;
  std::string param_string;
  int device = 0;
  int num_total_features = 0;
  std::vector<BinMapperH> mappers;       // per real feature
  std::vector<int> used;                 // inner -> real
  std::vector<int> real_to_inner;
  int max_num_bin = 1;
  int64_t num_data = 0;
  int64_t n_pad = 0;
  BinLayout layout{};
  std::vector<std::string> feature_names;

  // device-resident
  DevBuf<double> d_bounds;               // [F_used][kMaxBinsPerFeature], +inf padded
  DevBuf<FeatureMetaDev> d_fmeta;        // [F_used]
  DevBuf<int> d_real_index;              // [F_used]
  DevBuf<uint8_t> d_cm, d_rm;
  DevBuf<uint8_t> d_catbin;              // [F_used][kMaxBinsPerFeature]: categorical features — bin of the k-th smallest kept category
  bool has_categorical = false;
  DevBuf<float> d_label, d_weight;
  DevBuf<double> d_init_score;
  DevBuf<int32_t> d_query_boundaries;

  // host copies of the metadata (the caller frees its buffers right after each push; GetField returns pointers into these)
  std::vector<float> label, weight;
  std::vector<double> init_score;
  std::vector<int32_t> query_id, query_boundaries;
  bool has_weight = false, has_init_score = false, has_query = false;
  bool streaming_inited = false, finished = false, wait_for_manual_finish = false;
  int64_t pushed_rows = 0;

  ~Dataset();

  static Dataset* CreateFromSampledColumn(double** sample_data, int** sample_indices, int32_t ncol, const int* num_per_col,
                                          int32_t num_sample_row, int32_t num_local_row, int64_t num_dist_row,
                                          const char* params);
  static Dataset* CreateFromSerializedReference(const void* buf, int32_t len, int64_t num_row, const char* params);
  Dataset* CreateByReference(int64_t num_row) const;
  std::string SerializeReference() const;

  void InitStreaming(bool has_w, bool has_init, bool has_q, int nclasses, int nthreads);
  int num_init_score_classes = 1;       // init scores per row (multiclass: class-major [nclasses][num_data])
  float TimeBinning(const double* rows, int nrow, int repeats);      // measurement hook: ms per pass of the K2 kernel over nrow resident rows
  void PushRows(const void* data, int data_type, int32_t nrow, int32_t ncol, int64_t start_row, const float* lab,
                const float* w, const double* init, const int32_t* q, int tid);
  void PushRowsByCSR(const void* indptr, int indptr_type, const int32_t* indices, const void* data, int data_type, int64_t nindptr,
                     int64_t nelem, int64_t start_row, const float* lab, const float* w, const double* init, const int32_t* q, int tid);
  void PushMetadata(int32_t nrow, int64_t start_row, const float* lab, const float* w, const double* init, const int32_t* q);
  void MarkFinished();
  void SetFeatureNames(const char** names, int n);
  int num_features() const { return static_cast<int>(used.size()); }

  // introspection for parity tests
  void CopyBinColumn(int inner_feature, uint8_t* out) const;
  void CopyRowMajor(int64_t row, int64_t nrows, uint8_t* out) const;

 private:
  struct PushSlot {
    double* pinned[2] = {nullptr, nullptr};
    double* dstage[2] = {nullptr, nullptr};
    cudaEvent_t ev[2] = {nullptr, nullptr};
    cudaStream_t stream = nullptr;
    size_t cap_rows = 0;
    size_t pinned_bytes = 0;
    int cur = 0;
    int64_t seg_start = 0;
    int64_t seg_rows = 0;
  };
  std::deque<PushSlot> slots_;
  std::mutex mu_;
  void FinishSchema();
  void AllocateRows(int64_t n);
  void EnsureSlots(int n);
  void FlushSlot(PushSlot* s);
  void LaunchBinRows(const double* staged, int nrow, long long start_row, cudaStream_t stream);
  void FreeSlots();
};

int ResolveDevice(const Config& cfg);

}  // namespace lgb
