// Booster: objective kernels (K3), histogram construction (K4), sibling subtraction + split search (K5/K6), stable row
// partition (K7), score update (K8) and the leaf-wise / GBDT drivers.  Everything that happens inside lightgbmlib
// behind LGBM_BoosterCreate (lightgbm/.../booster/LightGBMBooster.scala:241, via TrainUtils.scala:16-29) and
// LGBM_BoosterUpdateOneIter (LightGBMBooster.scala:355-365, driven by TrainUtils.scala:98-135) lives here.
#pragma once
#include <deque>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "dataset.h"
#include "model.h"
#include "network.h"

namespace lgb {
struct NodeDev;
struct BatchCtrl;

struct SplitDev {
  double gain;
  int feature;          // inner index
  int real_feature;
  int threshold;
  int default_left;
  int left_count, right_count;
  double left_output, right_output;
  double left_sum_g, left_sum_h, right_sum_g, right_sum_h;
  long long left_int_g, left_int_h, right_int_g, right_int_h;
  int missing_type, nan_bin;   // of the split feature (nan_bin = -1: none), so that later stages need no feature table lookup
  // categorical split: the rows whose bin has its bit set in cat_bits go left, everything else (bin 0 included) right; threshold,
  // default_left and nan_bin are unused
  int is_cat;
  int monotone_type;           // direction of the split feature (monotone constraints; 0 for categorical splits)
  unsigned int cat_bits[8];
};

// ---- peer-memory exchange of the data-parallel learner (kernels: batch_tree.cuh) ------------------------------------------
// All ranks of one NVLink / NVSwitch node map each other's histogram work areas and one PeerMail block (CUDA IPC).  Signals
// are PUSHED: a rank stores its word into the mailbox of every peer (a remote store over NVLink) and every rank polls its own,
// local mailbox, so waiting costs no NVLink round trips.
constexpr int kMaxPeers = 8;
constexpr int kMaxBatch = 8;          // leaves expanded per round (batch_tree.cuh)
// The best split a rank found among ITS features for one child of the round (feature-sliced exchange), stamped with the exchange
// it belongs to.  tot_*: the leaf totals the rank's search saw (the root's come out of the search).
struct SplitMail {
  SplitDev s;
  long long tot_g, tot_h;
  unsigned int seq;
  unsigned int pad_[3];
};
struct PeerMail {
  unsigned int hist_flag[kMaxPeers];            // [r]: rank r's histograms of exchange #value are complete
  unsigned int max_seq[kMaxPeers];              // [r]: rank r's (max |g|, max |h|) of gradient pass #value has arrived
  unsigned int max_vals[2][kMaxPeers][2];       // [parity of the pass][r] = float bits of (max |g|, max |h|)
  unsigned int abort[kMaxPeers];                // [r] != 0: rank r gave up (error or time-out); everyone stops waiting
  unsigned int hist_count;                      // local: CTAs of the running histogram kernel that are finished
  unsigned int pad_[15];
  SplitMail split[2][2 * kMaxBatch][kMaxPeers]; // [parity of the exchange][batch slot * 2 + child][r]
};
// Passed to kernels by value.
struct PeerTable {
  const long long* work[kMaxPeers];     // every rank's work allocation (both parities), mapped into this process
  PeerMail* mail[kMaxPeers];            // every rank's mailbox; mail[rank] is this rank's own (local memory)
  int world;                            // 0: no exchange (single GPU, or the NCCL fallback)
  int rank;
  unsigned int seq;                     // the exchange this launch belongs to
  int sliced;                           // 1: rank g reduces and searches only features [f_lo, f_hi) and the best splits are exchanged
  int f_lo, f_hi;
  long long parity_off;                 // element offset of this exchange's buffer inside a rank's work allocation
  unsigned long long timeout_ns;        // a wait for a peer that lasts longer aborts the fit (listen_time_out of LGBM_NetworkInit)
};

struct LeafDev {
  int begin, count;      // local range of row indices
  int buf;               // index buffer holding the range; -1: identity (the root before its first split)
  int global_count;      // count the split finder uses: true count (serial) or the global estimate (data-parallel)
  long long int_g, int_h;
  double sum_g, sum_h;
  int depth, hist_slot;
  SplitDev best;
};

// Descriptors of the stand-alone histogram kernel k_hist (hist_kernel.cuh), used by the allocation-free entry point.
struct TreeCtrl {
  int done, find, smaller;
  int pad_;
  long long hist_rows, hist_launches;
};

// how a finished tree enters the scores (k_bt_apply_tree, batch_tree.cuh)
struct ScoreMix { double pre, post, bias; int shrink; };

struct TreeArraysDev {
  int *left_child, *right_child, *split_feature_inner, *threshold_in_bin, *leaf_count, *internal_count, *leaf_parent, *leaf_depth;
  signed char* decision_type;
  float* split_gain;
  double *leaf_value, *leaf_weight, *internal_value, *internal_weight;
  unsigned int* cat_bits;      // [L][8]: the bins that go left at a categorical node (nullptr: the dataset has no categorical feature)
};

struct LearnerParams {
  double l1, l2, max_delta_step, min_gain_to_split, min_sum_hessian, learning_rate;
  int min_data_in_leaf, max_depth, num_leaves, num_features;
  int quantized, dp, num_grad_quant_bins, stochastic_rounding, const_hessian;
  unsigned int seed;
  int reduce_in_select;                   // few features: the replay kernel reduces the per-feature results of a child (no last-arriver
                                          // reduction at the end of the split search)
  const unsigned char* feature_mask;      // [F] 1 = this tree may split on the feature (feature_fraction < 1), nullptr = all
  // categorical split search (upstream config.h: max_cat_to_onehot, max_cat_threshold, min_data_per_group, cat_smooth, cat_l2)
  int max_cat_to_onehot, max_cat_threshold, min_data_per_group;
  double cat_smooth, cat_l2;
  // monotone constraints, method "basic": per column +1 / -1 / 0, nullptr when no column is constrained (then nothing changes:
  // upstream's USE_MC code path — gains taken at clamped outputs — is on for EVERY feature as soon as one column is constrained)
  const signed char* monotone;
  double monotone_penalty;
  // feature_fraction_bynode < 1: [1 + 2 (L - 1)][F] column samples of this tree in the order the searched leaves use them
  // (nullptr otherwise); leaves are then expanded one per round, so that order is upstream's
  const unsigned char* node_mask;
};

class Booster {
 public:
  Booster(Dataset* train, const char* params);
  // scoring-only booster rebuilt from a model string (LGBM_BoosterLoadModelFromString); touches the GPU at the first prediction
  explicit Booster(const std::string& model_str);
  ~Booster();
  bool UpdateOneIter();
  bool UpdateOneIterCustom(const float* grad, const float* hess);     // host arrays of num_data floats each
  void ResetParameter(const char* params);
  std::string SaveModelToString(int start_iteration, int num_iteration) const;
  std::string DumpModel(int start_iteration, int num_iteration, int importance_type) const;      // JSON (LGBM_BoosterDumpModel)
  int NumTrees() const { return static_cast<int>(models_.size()); }
  int NumClasses() const { return K_; }
  int CurrentIteration() const { return iter_ + num_init_iteration_; }
  // LGBM_BoosterMerge (upstream GBDT::MergeFrom): the other booster's trees go in front of this one's; scores are not touched
  void MergeFrom(const Booster& other);
  void CopyScores(double* out) const;
  // scores of validation set i (LGBM_BoosterGetPredict with data_idx = i + 1), class-major; returns its row count
  int64_t CopyValidScores(int i, double* out) const;
  int64_t NumValidRows(int i) const;
  void CopyGradients(float* g, float* h) const;
  void CopyQuantGradients(int8_t* g, int8_t* h, double* scales) const;
  // GPU prediction on raw feature values (predict_kernel.cuh).  data_type: 0 float32, 1 float64; predict_type: 0 normal, 1 raw,
  // 2 leaf index.  Trees [start_iteration, start_iteration + num_iteration) (num_iteration <= 0: to the end).
  void PredictForMat(const void* data, int data_type, int64_t nrow, int ncol, bool row_major, int predict_type, int start_iteration,
                     int num_iteration, double* out, int64_t* out_len);
  int64_t CalcNumPredict(int64_t nrow, int predict_type, int start_iteration, int num_iteration) const;
  int NumTotalFeatures() const;
  // importance_type 0: number of splits per feature, 1: total split gain per feature; first num_iteration trees (<= 0: all)
  void FeatureImportance(int num_iteration, int importance_type, double* out) const;
  bool predict_only() const { return predict_only_; }
  // validation data and metrics (booster_eval.cu): data_idx 0 = training data, i >= 1 = the i-th validation set
  void AddValidData(Dataset* valid);
  std::vector<std::string> EvalNames();
  void Eval(int data_idx, std::vector<double>* out);
  const Dataset* train() const { return ds_; }

  // Kernel-level hooks used by parity tests, bench.py's roofline measurement and the profiler (no host sync inside):
  // build the histogram of the whole dataset (root) with the current quantized / fixed-point gradients into out
  // (F_used * 256 * 2 int64).
  void DebugRootHistogram(long long* out);
  float TimeRootHistogram(int repeats);
  int64_t last_hist_rows_visited() const { return last_rows_visited_; }
  int64_t hist_launches() const { return hist_launches_; }
  int64_t growth_rounds() const { return rounds_total_; }
  int64_t growth_expansions() const { return expansions_total_; }
  int64_t growth_splits() const { return splits_total_; }
  int64_t hist_kernel_launches() const { return hist_kernel_launches_; }
  double last_hist_ms() const { return last_hist_ms_; }
  int launches() const { return launches_; }
  void EnableHistTiming(bool on) { time_hist_ = on; }
  // per-kernel-type CUDA-event timing (slower: breaks launch overlap) and the set-up timings, as one JSON object
  void EnableProfile(bool on);
  std::string ProfileJson();
  // tells the other ranks of a data-parallel fit that this one gave up, so that they stop waiting (c_api.cu, on an exception)
  void AbortPeers();
  void TimerStart();
  float TimerStop();      // ms between TimerStart and now, measured with CUDA events on the launching stream

 private:
  Dataset* ds_;
  Config cfg_;
  LearnerParams lp_{};
  int device_ = 0;
  cudaStream_t stream_ = nullptr, copy_stream_ = nullptr;
  cudaEvent_t ev_rounds_done_ = nullptr;
  Network* net_ = nullptr;
  int64_t n_ = 0, n_global_ = 0, row_offset_ = 0;
  int F_ = 0;
  int iter_ = 0;
  enum Obj { kBinary, kL2, kLambdarank, kCustom, kHuber, kFair, kPoisson, kGamma, kTweedie, kMulticlass, kL1, kQuantile, kMape, kMulticlassOva, kXendcg, kXentropy } obj_ = kL2;
  DevBuf<unsigned int> d_xe_rand_;      // rank_xendcg: the LCG state of every query's stream
  bool PointRegression() const { return obj_ == kL2 || (obj_ >= kHuber && obj_ <= kTweedie) || RenewsLeaves(); }
  // l1 / quantile / mape (upstream IsRenewTreeOutput): every leaf's output is re-fit as a percentile of the residuals of its rows
  bool RenewsLeaves() const { return obj_ == kL1 || obj_ == kQuantile || obj_ == kMape; }
  // multiclassova (upstream MulticlassOVA: one BinaryLogloss per class, positive = "label == class"): per class the label
  // weights [neg, pos] and the (weighted) sums behind BoostFromScore, over all machines
  std::vector<double> ova_label_weights_, ova_suml_;
  double ova_sumw_ = 0.0;
  void RenewTreeOutput(HostTree* tree);
  double PercentileInitScore() const;
  std::vector<float> label_weight_;          // mape: 1 / max(1, |label|) [* weight]
  DevBuf<float> d_label_weight_, d_renew_w_[2];
  DevBuf<double> d_renew_res_[2], d_renew_out_;
  DevBuf<int> d_renew_off_;
  // boosting type (upstream src/boosting/{gbdt,rf,dart,goss}.hpp)
  enum Boost { kGBDT, kRF, kDART, kGOSS } boost_ = kGBDT;
  double shrinkage_rate_ = 0.1;             // learning_rate; DART rescales it by the number of dropped trees, RF uses 1
  ScoreMix mix_{1.0, 1.0, 0.0, 1};          // how the finished tree enters the scores (batch_tree.cuh)
  std::vector<double> rf_init_scores_;      // RF: the automatic initial scores; every tree carries them as a bias
  HostRandom drop_rand_{4};                 // DART: random_for_drop_
  std::vector<double> tree_weight_;
  double sum_weight_ = 0.0;
  std::vector<int> drop_index_;
  void DroppingTrees();
  void NormalizeDart();
  void GossSample();
  // adds an earlier tree (host copy, current leaf values) to the scores of class cls: DART takes trees out and puts them back
  void AddHostTreeToScores(const HostTree& t, int cls, bool train, bool valid);
  DevBuf<unsigned char> d_tree_blob2_;
  TreeArraysDev tree_dev2_{};
  std::vector<unsigned char> h_tree_blob2_;
  // multiclass: K_ trees per iteration, one per class; scores / gradients are class-major [K_][n_]; cls_ = the class whose tree is being trained
  int K_ = 1, cls_ = 0;
  std::vector<double> class_init_probs_;
  std::vector<char> class_need_train_;
  bool LogLink() const { return obj_ == kPoisson || obj_ == kGamma || obj_ == kTweedie; }      // predictions are exp(raw score)
  // kCustom: gradients come from the caller (objective=custom); kHuber.. : upstream regression_objective.hpp
  bool need_train_ = true, has_weights_ = false;
  double label_weights_[2] = {1.0, 1.0};
  std::vector<HostTree> models_;
  int num_init_iteration_ = 0;           // trees merged in front (continued training)
  // scoring: the model flattened into device arrays (re-uploaded when the model changed)
  bool predict_only_ = false;
  ModelHeader loaded_header_;
  long long model_version_ = 0, uploaded_version_ = -1;
  DevBuf<int> dp_node_off_, dp_leaf_off_, dp_feat_, dp_left_, dp_right_, dp_icount_, dp_lcount_, dp_depth_;
  DevBuf<int> dp_cat_begin_, dp_cat_words_;      // categorical nodes: first word / number of words of the node's category set
  DevBuf<unsigned int> dp_cat_thr_;
  DevBuf<double> dp_expected_;
  DevBuf<unsigned char> dp_paths_;         // feature contributions: the threads' TreeSHAP paths
  int model_max_depth_ = 0;
  DevBuf<double> dp_thr_, dp_leaf_, dp_out_;
  DevBuf<signed char> dp_dt_;
  DevBuf<unsigned char> dp_in_;
  void UploadModel();
  // Scoring entry points are called from many threads on one handle (LightGBMBooster.scala:67-140 keeps per-thread output
  // buffers for exactly that); the device staging buffers and the stream are per booster, so calls take turns.
  std::mutex predict_mu_;
  struct ValidSet {
    Dataset* ds = nullptr;
    DevBuf<double> score;
    DevBuf<double> inv_max_dcg, qweight;     // ndcg: [nq][K], [nq]
    bool ndcg_ready = false;
  };
  std::deque<ValidSet> valids_;
  ValidSet train_eval_;                      // metric constants of the training data (its scores are d_score_)
  struct MetricSpec { int kind; std::string name; };   // kind: 0 logloss, 1 error, 2 l2, 3 rmse, 4 auc, 5 ndcg (one name per eval_at)
  std::vector<MetricSpec> metrics_;
  bool metrics_ready_ = false;
  void InitMetrics();
  void PrepareNdcg(ValidSet* v, const Dataset* ds);
  void ApplyTreeToValidSets();
  void AddConstantToValidSets(double v);
  DevBuf<double2> d_mpart_;
  DevBuf<double> d_mout_, d_auc_keys_[2], d_auc_pos_, d_auc_neg_, d_auc_prefix_, d_ndcg_;
  DevBuf<int> d_auc_idx_[2];
  DevBuf<unsigned char> d_cub_tmp_;
  void TreeRange(int start_iteration, int num_iteration, int* t0, int* t1) const;

  // device state
  DevBuf<double> d_score_;
  // two gradient buffers: the fused apply-tree kernel writes the NEXT iteration's gradients while [grad_cur_] still holds the
  // ones the finished iteration trained on
  DevBuf<float> d_grad_[2], d_hess_[2];
  int grad_cur_ = 0;
  bool grad_ready_ = false;     // [grad_cur_ ^ 1] and the control block's max |g|, max |h| are valid for the next iteration
  bool fuse_grad_ = true;
  DevBuf<uint16_t> d_gh16_;          // quantized: (int8 grad << 8) | uint8 hess per row
  DevBuf<int2> d_ghfix_;             // float mode: fixed-point (grad, hess) per row
  DevBuf<int> d_idx_[2];
  DevBuf<long long> d_hist_pool_, d_hist_work_;
  NodeDev* d_nodes_ = nullptr;       // [max_nodes_]
  BatchCtrl* d_ctrl_ = nullptr;
  DevBuf<int> d_avail_;              // leaf id of the committed tree -> node
  int max_nodes_ = 0, spec_budget_ = 0, max_batch_ = 4, last_rounds_ = 0, round_slack_ = 1;
  long long rounds_enqueued_ = 0, refetches_ = 0;
  DevBuf<SplitDev> d_feat_best_;     // [2][F]
  DevBuf<unsigned char> d_splittable_;   // [num_leaves][F]
  DevBuf<unsigned char> d_feature_mask_; // [F] this tree's column sample (feature_fraction < 1 only)
  DevBuf<signed char> d_monotone_;       // [F] monotone direction per column (only with monotone_constraints)
  DevBuf<unsigned char> d_node_mask_;    // feature_fraction_bynode: this tree's by-node column samples
  HostRandom node_rand_saved_{2};        // the column sampler's stream before this tree's by-node samples were drawn
  int node_pool_size_ = 0, node_sample_cnt_ = 0, last_node_draws_ = 0, last_searched_ = 0;
  void SampleNodeMasksForTree();
  std::vector<unsigned char> tree_mask_host_;      // this tree's by-tree column sample (host copy)
  HostRandom feature_rand_{2};           // ColSampler's stream, seeded with feature_fraction_seed
  void SampleFeaturesForTree();
  bool TrainOneIter(const float* grad, const float* hess);
  // bagging (bagging_freq > 0, bagging_fraction < 1): the bag as a bit per row, re-drawn every bagging_freq iterations
  bool bagging_ = false;
  bool balanced_bagging_ = false;      // pos_bagging_fraction / neg_bagging_fraction (binary objective)
  DevBuf<unsigned int> d_bag_bits_, d_bag_rand_;
  DevBuf<int> d_bag_count_;
  long long bag_cnt_global_ = 0;
  void Bagging();
  DevBuf<unsigned char> d_tree_blob_;
  TreeArraysDev tree_dev_{};
  size_t tree_blob_bytes_ = 0;
  size_t tree_cat_off_ = 0;          // offset of the [L][8] category-set words in the tree blob (present with categorical features)
  std::vector<unsigned char> h_tree_blob_;
  // lambdarank
  DevBuf<double> d_sigmoid_table_, d_inv_max_dcg_, d_label_gain_, d_discount_;
  double min_sigmoid_input_ = -50, max_sigmoid_input_ = 50, sigmoid_table_idx_factor_ = 1;
  int max_query_len_ = 0;
  int short_query_len_ = 256, n_long_queries_ = 0;      // lambdarank launches are split by query length
  DevBuf<int> d_query_order_;

  // measurement
  bool time_hist_ = false;
  int64_t last_rows_visited_ = 0, hist_launches_ = 0;
  int64_t rounds_total_ = 0, expansions_total_ = 0, splits_total_ = 0, hist_kernel_launches_ = 0;   // batched growth statistics over all trees
  double last_hist_ms_ = 0;
  int launches_ = 0;
  int work_elems_ = 0;
  DevBuf<long long> d_trace_;
  int trace_cap_ = 0, trace_from_ = 0;
  // peer-memory histogram exchange (data-parallel learner)
  PeerTable peers_{};
  std::vector<void*> peer_work_, peer_mail_;
  DevBuf<PeerMail> d_peer_mail_;
  unsigned int exchange_seq_ = 0, max_seq_ = 0;
  bool nccl_exchange_ = false;       // peer memory could not be mapped: one NCCL all-reduce per round instead
  void SetupExchange();
  void CheckAbort(const BatchCtrl& hc);
  long long work_stride_ = 0;
  int part_blocks_ = 888;                   // 6 x 148 SMs (k_bt_partition is built for 6 CTAs per SM)
  int idx_es_ = 2;                  // ints per index entry (2: quantized, 4: float mode)
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> hist_events_;
  size_t hist_events_total_ = 0;      // events used by the running iteration (all of its trees)
  cudaEvent_t timer_ev_[2] = {nullptr, nullptr};
  // per-kernel-type timing with CUDA events (LGB_B200_PROFILE=1): development aid, breaks launch overlap
  bool prof_ = false;
  struct ProfRec { const char* name; cudaEvent_t a, b; };
  std::vector<ProfRec> prof_recs_;
  void ProfBegin(const char* name);
  void ProfEnd();
  void ProfReport();
  void ProfCollect();
  std::vector<std::pair<std::string, std::pair<long long, double>>> prof_agg_;      // name -> (launches, ms)
  std::vector<std::pair<std::string, double>> setup_ms_;                             // booster set-up phases
  long long part_rows_total_ = 0;                                                    // rows that went through the partition kernel

  void InitObjective();
  double BoostFromScoreLocalSums(double* suml, double* sumw) const;
  void ComputeGradients(const double* score_override = nullptr);
  void PrepareGradients(bool absmax_ready);           // absmax -> scales -> quantize / fixed point
  void TrainTree(HostTree* out);
  void LaunchHist(bool root);
  bool find_warp_ = false;      // split search: a warp (wide data) or a 256-thread CTA (narrow data) per (feature, child, batch item)
  void LaunchFind(bool root);
  void EnqueueRounds(int rounds);
  void LaunchSelect(bool after_root);
  void DumpTrace();
  bool FetchTree(HostTree* out);
  void UpdateScores();
  ModelHeader MakeHeader() const;
  void Sync() const;
};

}  // namespace lgb
