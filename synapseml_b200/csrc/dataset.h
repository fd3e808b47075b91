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

constexpr int kMaxBinsPerFeature = 256;   // bins of one uint8 column of the bin matrix
// A feature with more than 256 bins ("wide": max_bin > 255, what upstream stores as uint16) occupies S = ceil(num_bin / 253)
// consecutive uint8 columns.  Column c holds the logical bins [253 c, 253 c + 253) as the bytes 1..253; a row whose bin lives in a
// later column reads kWideAbove there, in an earlier column kWideBelow, and the rows of the feature's missing bin (NaN bin, or the
// zero bin under zero_as_missing) read kWideMissing in every column but their own.  With that code the histogram, partition and
// score kernels work on such a column unchanged: a split at logical bin t is the split (column t / 253, byte t % 253 + 1), where
// "byte <= threshold" sends the rows below left and the rows above right, and the missing byte follows default_left.  Only the
// row binning (writes S bytes per value) and the split search (scans the S columns as one feature) know about it.
constexpr int kWideLocalBins = 253;
constexpr int kWideBelow = 0, kWideMissing = 254, kWideAbove = 255;
constexpr int kMaxWideBins = 65535;

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
  // wide features (more than 256 bins): wide_cols = S in the feature's first column (whose num_bin / default_bin / most_freq_bin
  // are the LOGICAL ones), -k in its k-th further column, 0 for an ordinary feature; nan_local: the byte whose rows follow
  // default_left in this column (the missing bin's own byte in its home column, kWideMissing elsewhere, 256 = no missing bin);
  // wide_off: the feature's upper bounds in Dataset::d_wide_bounds
  int wide_cols, nan_local, wide_off;
  int pad_;
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
  std::vector<int> used;                 // inner column -> real feature (a wide feature has several consecutive columns)
  std::vector<int> real_to_inner;        // real feature -> its first inner column (-1: trivial)
  std::vector<int> logical_used;         // the non-trivial real features in order: what upstream calls the inner features
  std::vector<int> logical_first_col, logical_num_cols;      // per logical feature
  bool has_wide = false;
  DevBuf<double> d_wide_bounds;          // upper bounds of the wide features, feature after feature
  int num_logical_features() const { return static_cast<int>(logical_used.size()); }
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
                                          const char* params, const double* dense_rows = nullptr);      // dense_rows: the sample as a row-major matrix instead (LGBM_B200_DatasetCreateFromSampledRows)
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
  void CopyBinColumn(int logical_feature, uint16_t* out) const;      // bins of the k-th non-trivial feature (wide features decoded)
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
