#include "booster.h"

#include <cub/device/device_segmented_radix_sort.cuh>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstring>
#include <map>

#include "batch_tree.cuh"
#include "kernels.cuh"
#include "predict_kernel.cuh"

namespace lgb {

std::string FormatG17(double v);
std::string FormatG(double v);

namespace {
constexpr int kNumSms = 148;

template <typename K>
void SetSmem(K kernel, size_t bytes) {
  LGB_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(bytes)));
}
// Launch with programmatic stream serialization: the kernel may become resident while its predecessor drains; it calls
// PdlWait() before touching anything the predecessor wrote.
template <typename... KArgs, typename... Args>
void LaunchPdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, Args... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  LGB_CUDA(cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...));
}
constexpr size_t kHistSmem = kHistSmemBytes;   // 4 x 32 KB arrays + 32 KB alignment slack, both modes
}  // namespace

Booster::Booster(Dataset* train, const char* params) : ds_(train) {
  LGB_CHECK(train != nullptr, "BoosterCreate: training dataset is null");
  LGB_CHECK(train->finished, "BoosterCreate: the training dataset is not finished (LGBM_DatasetMarkFinished was not called)");
  LGB_CHECK(train->num_data > 0, "BoosterCreate: the training dataset has no rows");
  cfg_ = train->cfg;
  cfg_.raw.clear();
  cfg_.Set(params ? params : "");
  // dataset-time parameters the booster string does not repeat keep their dataset values
  for (auto& kv : train->cfg.raw)
    if (!cfg_.raw.count(kv.first)) cfg_.raw[kv.first] = kv.second;
  g_verbosity = cfg_.verbosity;
  LGB_CHECK(cfg_.boosting == "gbdt" || cfg_.boosting == "rf" || cfg_.boosting == "dart" || cfg_.boosting == "goss",
            "boosting_type=%s is not supported by this library (gbdt, rf, dart, goss)", cfg_.boosting.c_str());
  boost_ = cfg_.boosting == "rf" ? kRF : cfg_.boosting == "dart" ? kDART : cfg_.boosting == "goss" ? kGOSS : kGBDT;
  shrinkage_rate_ = boost_ == kRF ? 1.0 : cfg_.learning_rate;
  drop_rand_ = HostRandom(cfg_.drop_seed);
  if (boost_ == kRF)
    LGB_CHECK((cfg_.bagging_freq > 0 && cfg_.bagging_fraction < 1.0) || cfg_.feature_fraction < 1.0,
              "boosting=rf needs bagging (bagging_freq > 0 and bagging_fraction < 1) or feature_fraction < 1");
  if (boost_ == kGOSS) {
    LGB_CHECK(cfg_.top_rate > 0.0 && cfg_.other_rate > 0.0 && cfg_.top_rate + cfg_.other_rate <= 1.0, "goss: top_rate + other_rate must be in (0, 1]");
    LGB_CHECK(!(cfg_.bagging_freq > 0 && cfg_.bagging_fraction < 1.0), "goss cannot be combined with bagging");
  }
  LGB_CHECK(cfg_.tree_learner == "serial" || cfg_.tree_learner == "data" || cfg_.tree_learner == "data_parallel",
            "tree_learner=%s is not supported by this library (data_parallel / serial)", cfg_.tree_learner.c_str());
  LGB_CHECK(cfg_.num_leaves >= 2 && cfg_.num_leaves <= 1024, "num_leaves=%d is outside [2, 1024]", cfg_.num_leaves);
  LGB_CHECK(train->num_features() <= kApplyMaxFeatures, "more than %d used features are not supported", kApplyMaxFeatures);
  LGB_CHECK(cfg_.num_grad_quant_bins >= 2 && cfg_.num_grad_quant_bins <= 64, "num_grad_quant_bins=%d is outside [2, 64]", cfg_.num_grad_quant_bins);
  LGB_CHECK(cfg_.feature_fraction > 0.0 && cfg_.feature_fraction <= 1.0, "feature_fraction=%g is outside (0, 1]", cfg_.feature_fraction);
  // parameters that would change the model and are not implemented are refused rather than ignored
  LGB_CHECK(cfg_.feature_fraction_bynode > 0.0 && cfg_.feature_fraction_bynode <= 1.0, "feature_fraction_bynode=%g is outside (0, 1]",
            cfg_.feature_fraction_bynode);
  LGB_CHECK(cfg_.bagging_fraction > 0.0 && cfg_.bagging_fraction <= 1.0, "bagging_fraction=%g is outside (0, 1]", cfg_.bagging_fraction);
  LGB_CHECK(cfg_.pos_bagging_fraction > 0.0 && cfg_.pos_bagging_fraction <= 1.0 && cfg_.neg_bagging_fraction > 0.0 && cfg_.neg_bagging_fraction <= 1.0,
            "pos_bagging_fraction / neg_bagging_fraction must be in (0, 1]");
  // (max_bin_by_feature belongs to dataset construction: given to the booster only, upstream warns that it cannot change a
  // constructed dataset and goes on, which is what the reference's flow does — LightGBMBase.scala:374-386 never sends it there)
  if (cfg_.raw.count("max_bin_by_feature") && train->cfg.max_bin_by_feature.empty())
    LogWarning("Cannot change max_bin_by_feature after constructed Dataset handle.");
  for (const char* key : {"interaction_constraints", "forcedsplits_filename",
                          "cegb_penalty_split", "linear_tree", "extra_trees", "path_smooth", "feature_contri"}) {
    auto it = cfg_.raw.find(key);
    if (it == cfg_.raw.end()) continue;
    const std::string& v = it->second;
    const bool off = v.empty() || v == "0" || v == "0.0" || v == "false" || v == "False";
    LGB_CHECK(off, "%s=%s is not supported by this library yet", key, v.c_str());
  }
  device_ = train->device;
  LGB_CUDA(cudaSetDevice(device_));
  LGB_CUDA(cudaStreamCreateWithFlags(&stream_, cudaStreamNonBlocking));
  LGB_CUDA(cudaStreamCreateWithFlags(&copy_stream_, cudaStreamNonBlocking));
  LGB_CUDA(cudaEventCreateWithFlags(&ev_rounds_done_, cudaEventDisableTiming));
  // Histogram passes over non-root leaves gather one 32-byte row record per visited row: ask L2 not to over-fetch 64 bytes
  // per miss (the default granularity), which doubles the DRAM traffic of sparse gathers.  A hint; failure is harmless.
  if (getenv("LGB_B200_L2_FETCH_64") == nullptr) cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, 32);
  net_ = Network::Get();
  n_ = train->num_data;
  F_ = train->num_features();
  has_weights_ = train->has_weight;
  // global row count and this shard's offset (rank order = row order, BasePartitionTask.scala:92-112)
  std::vector<long long> all;
  net_->HostAllGatherInt64(n_, &all);
  n_global_ = 0;
  row_offset_ = 0;
  for (int r = 0; r < net_->world(); ++r) {
    if (r < net_->rank()) row_offset_ += all[r];
    n_global_ += all[r];
  }
  LGB_CHECK(n_global_ < (int64_t(1) << 31), "more than 2^31 rows in total are not supported");
  InitObjective();
  lp_.l1 = cfg_.lambda_l1; lp_.l2 = cfg_.lambda_l2; lp_.max_delta_step = cfg_.max_delta_step;
  lp_.min_gain_to_split = cfg_.min_gain_to_split; lp_.min_sum_hessian = cfg_.min_sum_hessian_in_leaf;
  lp_.learning_rate = cfg_.learning_rate; lp_.min_data_in_leaf = cfg_.min_data_in_leaf; lp_.max_depth = cfg_.max_depth;
  lp_.num_leaves = cfg_.num_leaves; lp_.num_features = F_;
  lp_.quantized = cfg_.use_quantized_grad; lp_.dp = net_->active();
  lp_.num_grad_quant_bins = cfg_.num_grad_quant_bins; lp_.stochastic_rounding = cfg_.stochastic_rounding;
  lp_.const_hessian = ((obj_ == kL2 || obj_ == kL1 || obj_ == kQuantile) && !has_weights_) || obj_ == kMape;      // IsConstantHessian
  lp_.seed = static_cast<unsigned int>(cfg_.seed);
  lp_.max_cat_to_onehot = cfg_.max_cat_to_onehot; lp_.max_cat_threshold = cfg_.max_cat_threshold;
  lp_.min_data_per_group = cfg_.min_data_per_group; lp_.cat_smooth = cfg_.cat_smooth; lp_.cat_l2 = cfg_.cat_l2;
  // 32 features or fewer: a warp of the replay kernel reduces a child's per-feature results (LGB_B200_REDUCE_IN_SELECT=0: the split
  // search's last unit does, as for wide data)
  lp_.reduce_in_select = F_ <= 32 && !(getenv("LGB_B200_REDUCE_IN_SELECT") && atoi(getenv("LGB_B200_REDUCE_IN_SELECT")) == 0);
  // split search: a CTA per unit while a round's units fit about one wave of CTAs (8 per SM), a warp per unit beyond that
  // (with histograms exchanged over peer memory the CTA form stays: a unit's bins are read from every rank, and 256 threads with one
  // remote load per rank each hide NVLink latency better than a lane with eight — C5 on 8 GPUs: 0.94 vs 2.1 ms per iteration)
  find_warp_ = static_cast<long long>(F_) * 2 * kMaxBatch > 148LL * 8 * 2 && !net_->active();
  if (const char* env = getenv("LGB_B200_FIND_WARP")) find_warp_ = atoi(env) != 0;
  if (ds_->has_categorical || ds_->has_wide) find_warp_ = false;      // the categorical and the wide-feature search exist in the CTA form only
  // monotone constraints (LightGBMParams.scala monotoneConstraints / monotoneConstraintsMethod / monotonePenalty), method "basic"
  lp_.monotone = nullptr;
  lp_.monotone_penalty = cfg_.monotone_penalty;
  if (std::any_of(cfg_.monotone_constraints.begin(), cfg_.monotone_constraints.end(), [](int v) { return v != 0; })) {
    LGB_CHECK(cfg_.monotone_constraints_method == "basic", "monotone_constraints_method=%s is not supported by this library (basic)",
              cfg_.monotone_constraints_method.c_str());
    LGB_CHECK(static_cast<int>(cfg_.monotone_constraints.size()) == ds_->num_total_features,
              "monotone_constraints has %d values for %d columns", static_cast<int>(cfg_.monotone_constraints.size()), ds_->num_total_features);
    LGB_CHECK(cfg_.monotone_penalty >= 0.0, "monotone_penalty must not be negative");
    std::vector<signed char> mono(std::max(F_, 1), 0);
    for (int f = 0; f < F_; ++f) {
      const int v = cfg_.monotone_constraints[ds_->used[f]];
      LGB_CHECK(v >= -1 && v <= 1, "monotone_constraints values must be -1, 0 or 1");
      LGB_CHECK(v == 0 || ds_->mappers[ds_->used[f]].bin_type == 0, "a categorical feature cannot carry a monotone constraint");
      mono[f] = static_cast<signed char>(v);
    }
    d_monotone_.Alloc(mono.size());
    LGB_CUDA(cudaMemcpy(d_monotone_.p, mono.data(), mono.size(), cudaMemcpyHostToDevice));
    lp_.monotone = d_monotone_.p;
    find_warp_ = false;      // the constrained search exists in the CTA form only
  }
  lp_.feature_mask = nullptr;
  feature_rand_ = HostRandom(cfg_.feature_fraction_seed);
  if (cfg_.feature_fraction < 1.0 && F_ > 0) {
    d_feature_mask_.Alloc(F_);
    lp_.feature_mask = d_feature_mask_.p;
  }
  // feature_fraction_bynode (LightGBMParams.scala featureFractionByNode; ColSampler::GetByNode): every searched leaf draws its own
  // column sample from one stream, in upstream's sequential order — so leaves are expanded one per round (no speculation), and the
  // tree's samples are drawn ahead on the host in the order the device will use them
  lp_.node_mask = nullptr;
  if (cfg_.feature_fraction_bynode < 1.0 && F_ > 0) {
    d_node_mask_.Alloc(static_cast<size_t>(1 + 2 * (cfg_.num_leaves - 1)) * F_);
    lp_.node_mask = d_node_mask_.p;
    find_warp_ = false;
  }

  const int L = cfg_.num_leaves;
  d_score_.AllocZero(n_ * K_);      // class-major [K_][n_]
  if (train->has_init_score) {
    LGB_CHECK(static_cast<int64_t>(train->init_score.size()) == n_ * K_, "init_score holds %lld values, expected num_data * num_class = %lld",
              static_cast<long long>(train->init_score.size()), static_cast<long long>(n_ * K_));
    LGB_CUDA(cudaMemcpy(d_score_.p, train->d_init_score.p, sizeof(double) * n_ * K_, cudaMemcpyDeviceToDevice));
  }
  d_grad_[0].Alloc(n_ * K_); d_hess_[0].Alloc(n_ * K_);
  d_grad_[1].Alloc(K_ == 1 ? n_ : 1); d_hess_[1].Alloc(K_ == 1 ? n_ : 1);      // the second pair only serves the fused apply + gradient pass
  if (K_ > 1) fuse_grad_ = false;      // a row's gradients depend on all K scores: one gradient kernel per iteration
  if (boost_ == kRF || boost_ == kDART) fuse_grad_ = false;      // RF: gradients are taken once; DART: trees leave and re-enter the scores between iterations
  if (const char* env = getenv("LGB_B200_FUSE_GRAD")) fuse_grad_ = atoi(env) != 0 && K_ == 1;
  if (lp_.quantized) d_gh16_.Alloc(n_); else d_ghfix_.Alloc(n_);
  idx_es_ = lp_.quantized ? 2 : 4;      // ints per index entry: {row, g|h} or {row, g, h, -}
  d_idx_[0].Alloc(n_ * idx_es_); d_idx_[1].Alloc(n_ * idx_es_);
  // batched best-first learner (batch_tree.cuh): up to max_batch_ leaves are expanded per round; spec_budget_ bounds the
  // expansions that are not guaranteed to be committed
  {
    const char* env = getenv("LGB_B200_MAX_BATCH");
    max_batch_ = env ? std::max(1, std::min(kMaxBatch, atoi(env))) : kMaxBatch;
    if (lp_.node_mask != nullptr) max_batch_ = 1;
  }
  if (const char* env = getenv("LGB_B200_PART_BLOCKS")) part_blocks_ = std::max(1, atoi(env));
  if (const char* env = getenv("LGB_B200_ROUND_SLACK")) round_slack_ = atoi(env);      // experiments
  spec_budget_ = std::max(2 * kMaxBatch, (L - 1) / 2);
  if (const char* env = getenv("LGB_B200_SPEC_BUDGET")) spec_budget_ = std::max(0, atoi(env));      // experiments
  const int max_expansions = (L - 1) + spec_budget_;
  max_nodes_ = 1 + 2 * max_expansions;
  const size_t hist_elems = static_cast<size_t>(std::max(F_, 1)) * kBins * 2;
  d_hist_pool_.Alloc(hist_elems * (1 + max_expansions));
  d_hist_work_.Alloc(hist_elems * kMaxBatch * 2);      // two parities (only the first is used without the peer exchange)
  work_stride_ = static_cast<long long>(hist_elems) * kMaxBatch;
  SetupExchange();
  // quantized mode packs (sum g, sum h) of a bin into one int64 (hist_kernel.cuh), float mode keeps two
  work_elems_ = static_cast<int>(lp_.quantized ? hist_elems / 2 : hist_elems);
  LGB_CHECK(!lp_.quantized || static_cast<double>(n_global_) * cfg_.num_grad_quant_bins < 4294967296.0,
            "use_quantized_grad: num_data * num_grad_quant_bins must stay below 2^32");
  LGB_CUDA(cudaMalloc(&d_nodes_, sizeof(NodeDev) * max_nodes_));
  LGB_CUDA(cudaMalloc(&d_ctrl_, sizeof(BatchCtrl)));
  LGB_CUDA(cudaMemset(d_ctrl_, 0, sizeof(BatchCtrl)));
  if (getenv("LGB_B200_TRACE") != nullptr) {      // experiments only: per-kernel timeline of the tree kernels
    trace_cap_ = 1 << 16;
    d_trace_.AllocZero(static_cast<size_t>(2) * trace_cap_);
    BatchCtrl hc;
    memset(&hc, 0, sizeof(hc));
    hc.trace = nullptr; hc.trace_cap = trace_cap_;
    LGB_CUDA(cudaMemcpy(d_ctrl_, &hc, sizeof(hc), cudaMemcpyHostToDevice));
    if (const char* from = getenv("LGB_B200_TRACE_FROM")) trace_from_ = atoi(from);
  }
  // ResetSampleConfig: balanced bagging (pos / neg_bagging_fraction) needs a binary objective with positive rows on this machine
  balanced_bagging_ = cfg_.bagging_freq > 0 && (cfg_.pos_bagging_fraction < 1.0 || cfg_.neg_bagging_fraction < 1.0) && obj_ == kBinary &&
                      std::any_of(ds_->label.begin(), ds_->label.end(), [](float l) { return l > 0; });
  bagging_ = (cfg_.bagging_freq > 0 && (cfg_.bagging_fraction < 1.0 || balanced_bagging_)) || boost_ == kGOSS;
  if (bagging_) {      // ResetSampleConfig: one LCG stream per block of 1024 local rows, seeded bagging_seed + block
    const int nblk = static_cast<int>((n_ + 1023) / 1024);
    std::vector<unsigned int> seeds(nblk);
    for (int b = 0; b < nblk; ++b) seeds[b] = static_cast<unsigned int>(cfg_.bagging_seed + b);
    d_bag_rand_.Alloc(nblk);
    LGB_CUDA(cudaMemcpy(d_bag_rand_.p, seeds.data(), sizeof(unsigned int) * nblk, cudaMemcpyHostToDevice));
    d_bag_bits_.AllocZero(static_cast<size_t>(nblk) * 32);
    d_bag_count_.AllocZero(1);
    const unsigned int* bits = d_bag_bits_.p;
    LGB_CUDA(cudaMemcpy(&d_ctrl_->bag_bits, &bits, sizeof(bits), cudaMemcpyHostToDevice));
  }
  d_avail_.AllocZero(L);
  d_feat_best_.Alloc(static_cast<size_t>(2) * kMaxBatch * std::max(F_, 1));
  d_splittable_.Alloc(static_cast<size_t>(max_nodes_) * std::max(F_, 1));
  // tree arrays blob: 8 int arrays [L], decision_type [L], split_gain [L], 4 double arrays [L]
  {
    const size_t ints = 8 * sizeof(int) * L, chars = RoundUp<size_t>(L, 8), floats = RoundUp<size_t>(sizeof(float) * L, 8), dbl = 4 * sizeof(double) * L;
    tree_cat_off_ = dbl + ints + floats + chars;      // categorical features: [L][8] words behind the other arrays
    tree_blob_bytes_ = tree_cat_off_ + (ds_->has_categorical ? sizeof(unsigned int) * 8 * L : 0);
    d_tree_blob_.AllocZero(tree_blob_bytes_);
    h_tree_blob_.resize(tree_blob_bytes_);
    unsigned char* b = d_tree_blob_.p;
    tree_dev_.leaf_value = reinterpret_cast<double*>(b); tree_dev_.leaf_weight = tree_dev_.leaf_value + L;
    tree_dev_.internal_value = tree_dev_.leaf_weight + L; tree_dev_.internal_weight = tree_dev_.internal_value + L;
    int* ip = reinterpret_cast<int*>(b + dbl);
    tree_dev_.left_child = ip; tree_dev_.right_child = ip + L; tree_dev_.split_feature_inner = ip + 2 * L;
    tree_dev_.threshold_in_bin = ip + 3 * L; tree_dev_.leaf_count = ip + 4 * L; tree_dev_.internal_count = ip + 5 * L;
    tree_dev_.leaf_parent = ip + 6 * L; tree_dev_.leaf_depth = ip + 7 * L;
    tree_dev_.split_gain = reinterpret_cast<float*>(b + dbl + ints);
    tree_dev_.decision_type = reinterpret_cast<signed char*>(b + dbl + ints + floats);
    tree_dev_.cat_bits = ds_->has_categorical ? reinterpret_cast<unsigned int*>(b + tree_cat_off_) : nullptr;
  }
  if (boost_ == kDART) {      // a second set of tree arrays for the earlier trees DART re-adds to the scores
    d_tree_blob2_.AllocZero(tree_blob_bytes_);
    h_tree_blob2_.resize(tree_blob_bytes_);
    const size_t ints = 8 * sizeof(int) * L, floats = RoundUp<size_t>(sizeof(float) * L, 8), dbl = 4 * sizeof(double) * L;
    unsigned char* b = d_tree_blob2_.p;
    tree_dev2_.leaf_value = reinterpret_cast<double*>(b); tree_dev2_.leaf_weight = tree_dev2_.leaf_value + L;
    tree_dev2_.internal_value = tree_dev2_.leaf_weight + L; tree_dev2_.internal_weight = tree_dev2_.internal_value + L;
    int* ip = reinterpret_cast<int*>(b + dbl);
    tree_dev2_.left_child = ip; tree_dev2_.right_child = ip + L; tree_dev2_.split_feature_inner = ip + 2 * L;
    tree_dev2_.threshold_in_bin = ip + 3 * L; tree_dev2_.leaf_count = ip + 4 * L; tree_dev2_.internal_count = ip + 5 * L;
    tree_dev2_.leaf_parent = ip + 6 * L; tree_dev2_.leaf_depth = ip + 7 * L;
    tree_dev2_.split_gain = reinterpret_cast<float*>(b + dbl + ints);
    tree_dev2_.decision_type = reinterpret_cast<signed char*>(b + dbl + ints + floats);
    tree_dev2_.cat_bits = ds_->has_categorical ? reinterpret_cast<unsigned int*>(b + tree_cat_off_) : nullptr;
  }
  SetSmem(k_bt_select, SelectSmemBytes(1024));   // the largest tree the learner accepts; launches ask for what their num_leaves needs
#define SET_HIST_SMEM(RR, BPL, FLT) SetSmem(k_bt_hist<RR, BPL, FLT, true>, kHistSmem); SetSmem(k_bt_hist<RR, BPL, FLT, false>, kHistSmem)
  SET_HIST_SMEM(4, 4, false); SET_HIST_SMEM(2, 4, false); SET_HIST_SMEM(1, 4, false);
  SET_HIST_SMEM(4, 2, true); SET_HIST_SMEM(2, 2, true); SET_HIST_SMEM(1, 2, true);
#undef SET_HIST_SMEM
  LGB_CUDA(cudaStreamSynchronize(stream_));
  prof_ = getenv("LGB_B200_PROFILE") != nullptr;
}

// Data-parallel learner: how the ranks exchange histograms.  Peer memory (all ranks on one NVLink / NVSwitch node) needs no
// communicator; only if it cannot be mapped is NCCL loaded and initialised (seconds at 8 ranks, so not on the default path).
void Booster::SetupExchange() {
  memset(&peers_, 0, sizeof(peers_));
  if (!net_->active()) return;
  const auto t0 = std::chrono::steady_clock::now();
  const int world = net_->world();
  const size_t hist_elems = static_cast<size_t>(std::max(F_, 1)) * kBins * 2;
  const bool want_peer = world <= kMaxPeers && (getenv("LGB_B200_PEER_EXCHANGE") == nullptr || atoi(getenv("LGB_B200_PEER_EXCHANGE")) != 0);
  if (want_peer) {
    d_peer_mail_.AllocZero(1);
    LGB_CUDA(cudaDeviceSynchronize());      // the zeroed mailbox must be in place before any peer can write to it
    if (net_->ExchangePeerPointers(device_, d_hist_work_.p, &peer_work_) && net_->ExchangePeerPointers(device_, d_peer_mail_.p, &peer_mail_)) {
      for (int r = 0; r < world; ++r) {
        peers_.work[r] = static_cast<const long long*>(peer_work_[r]);
        peers_.mail[r] = static_cast<PeerMail*>(peer_mail_[r]);
      }
      peers_.world = world;
      peers_.rank = net_->rank();
      peers_.timeout_ns = static_cast<unsigned long long>(std::max(1, net_->timeout_s())) * 1000000000ULL;
      // Feature-sliced exchange for wide data.  All-to-all moves world x (the round's bins) into every rank and searches every
      // feature world times: right while a round is latency-bound (C2: 8 x 57 KB per rank), wrong for C5 (8 x 4 MB per rank).
      const double all_to_all_bytes = static_cast<double>(lp_.quantized ? hist_elems / 2 : hist_elems) * sizeof(long long) * kMaxBatch * world;
      bool sliced = F_ >= world && all_to_all_bytes > 16.0 * 1024 * 1024;
      if (const char* env = getenv("LGB_B200_SLICED")) sliced = atoi(env) != 0 && F_ >= world;
      if (ds_->has_wide) sliced = false;      // a wide feature's columns are searched by one unit: they must not straddle two ranks
      peers_.sliced = sliced ? 1 : 0;
      peers_.f_lo = sliced ? static_cast<int>(static_cast<long long>(F_) * peers_.rank / world) : 0;
      peers_.f_hi = sliced ? static_cast<int>(static_cast<long long>(F_) * (peers_.rank + 1) / world) : F_;
      LogInfo("Histogram exchange over peer memory (%d ranks, %s)", world, sliced ? "feature-sliced" : "all-to-all");
      if (getenv("LGB_B200_PEER_DEBUG"))
        fprintf(stderr, "[peer] rank %d: exchange over peer memory, %d ranks, %s\n", net_->rank(), world, sliced ? "feature-sliced" : "all-to-all");
    } else {
      LogWarning("Peer memory could not be mapped; histograms are exchanged with NCCL all-reduce");
      if (getenv("LGB_B200_PEER_DEBUG")) fprintf(stderr, "[peer] rank %d: mapping failed, NCCL fallback\n", net_->rank());
    }
  }
  if (peers_.world == 0) {
    net_->EnsureNccl(device_);
    nccl_exchange_ = true;
    if (getenv("LGB_B200_PEER_DEBUG")) fprintf(stderr, "[peer] rank %d: NCCL all-reduce exchange\n", net_->rank());
  }
  setup_ms_.emplace_back("exchange_setup", std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count());
}

Booster::~Booster() {
  if (device_ < 0) return;      // a loaded model that never predicted owns nothing on a GPU
  cudaSetDevice(device_);
  try { ProfReport(); } catch (...) {}
  try { DumpTrace(); } catch (...) {}
  if (peers_.world > 0) {
    // a peer may still be reading this rank's last histograms / writing its mailbox: leave together
    try { std::vector<long long> all; net_->HostAllGatherInt64(0, &all); } catch (...) {}
    net_->ClosePeerPointers(peer_work_); net_->ClosePeerPointers(peer_mail_);
  }
  if (d_nodes_) cudaFree(d_nodes_);
  if (d_ctrl_) cudaFree(d_ctrl_);
  if (stream_) { cudaStreamSynchronize(stream_); cudaStreamDestroy(stream_); }
  if (copy_stream_) cudaStreamDestroy(copy_stream_);
  if (ev_rounds_done_) cudaEventDestroy(ev_rounds_done_);
  for (auto& e : hist_events_) { cudaEventDestroy(e.first); cudaEventDestroy(e.second); }
  if (timer_ev_[0]) { cudaEventDestroy(timer_ev_[0]); cudaEventDestroy(timer_ev_[1]); }
}

void Booster::ProfBegin(const char* name) {
  if (!prof_) return;
  ProfRec r{name, nullptr, nullptr};
  LGB_CUDA(cudaEventCreate(&r.a)); LGB_CUDA(cudaEventCreate(&r.b));
  LGB_CUDA(cudaEventRecord(r.a, stream_));
  prof_recs_.push_back(r);
}
void Booster::ProfEnd() {
  if (!prof_) return;
  LGB_CUDA(cudaEventRecord(prof_recs_.back().b, stream_));
}
void Booster::ProfCollect() {
  if (prof_recs_.empty()) return;
  LGB_CUDA(cudaStreamSynchronize(stream_));
  for (auto& r : prof_recs_) {
    float ms = 0; cudaEventElapsedTime(&ms, r.a, r.b);
    auto it = std::find_if(prof_agg_.begin(), prof_agg_.end(), [&](const auto& e) { return e.first == r.name; });
    if (it == prof_agg_.end()) { prof_agg_.emplace_back(r.name, std::make_pair(0LL, 0.0)); it = prof_agg_.end() - 1; }
    it->second.first += 1; it->second.second += ms;
    cudaEventDestroy(r.a); cudaEventDestroy(r.b);
  }
  prof_recs_.clear();
}
void Booster::ProfReport() {
  if (!prof_ || getenv("LGB_B200_PROFILE") == nullptr) return;
  ProfCollect();
  for (auto& kv : prof_agg_)
    fprintf(stderr, "[prof] %-16s n=%6lld total=%9.3f ms avg=%8.2f us\n", kv.first.c_str(), kv.second.first, kv.second.second,
            1e3 * kv.second.second / std::max(1LL, kv.second.first));
}
void Booster::EnableProfile(bool on) {
  LGB_CUDA(cudaSetDevice(device_));
  ProfCollect();
  prof_ = on;
}
// {"setup_ms": {...}, "kernels": {"name": {"launches": n, "ms": t}, ...}, "partition_rows": r}
std::string Booster::ProfileJson() {
  LGB_CUDA(cudaSetDevice(device_));
  ProfCollect();
  std::string o = "{\"setup_ms\": {";
  for (size_t i = 0; i < setup_ms_.size(); ++i)
    o += StrFormat("%s\"%s\": %.3f", i ? ", " : "", setup_ms_[i].first.c_str(), setup_ms_[i].second);
  o += "}, \"kernels\": {";
  for (size_t i = 0; i < prof_agg_.size(); ++i)
    o += StrFormat("%s\"%s\": {\"launches\": %lld, \"ms\": %.6f}", i ? ", " : "", prof_agg_[i].first.c_str(), prof_agg_[i].second.first,
                   prof_agg_[i].second.second);
  o += StrFormat("}, \"partition_rows\": %lld, \"rounds_enqueued\": %lld, \"rounds_used\": %lld, \"tree_refetches\": %lld, \"exchange\": \"%s\"}",
                 part_rows_total_, rounds_enqueued_, static_cast<long long>(rounds_total_), refetches_,
                 peers_.world > 0 ? (peers_.sliced ? "peer-sliced" : "peer-all-to-all") : nccl_exchange_ ? "nccl" : "none");
  return o;
}

void Booster::DumpTrace() {
  if (trace_cap_ == 0) return;
  BatchCtrl hc;
  LGB_CUDA(cudaMemcpy(&hc, d_ctrl_, sizeof(hc), cudaMemcpyDeviceToHost));
  const int n = std::min(hc.trace_n, trace_cap_);
  std::vector<long long> t(static_cast<size_t>(2) * n);
  LGB_CUDA(cudaMemcpy(t.data(), d_trace_.p, sizeof(long long) * 2 * n, cudaMemcpyDeviceToHost));
  std::string path = getenv("LGB_B200_TRACE");
  if (net_ != nullptr && net_->active()) path += ".r" + std::to_string(net_->rank());      // one timeline per rank
  FILE* f = fopen(path.c_str(), "w");
  if (!f) return;
  for (int i = 0; i < n; ++i) fprintf(f, "%lld %lld\n", t[2 * i], t[2 * i + 1]);
  fclose(f);
}

void Booster::Sync() const { LGB_CUDA(cudaStreamSynchronize(stream_)); }

void Booster::ResetParameter(const char* params) {
  Config c;
  c.Set(params ? params : "");
  if (c.raw.count("learning_rate")) {
    LGB_CHECK(c.learning_rate > 0.0, "learning_rate must be positive");
    cfg_.learning_rate = c.learning_rate; lp_.learning_rate = c.learning_rate; cfg_.raw["learning_rate"] = c.raw["learning_rate"];
    // GBDT::ResetConfig: the shrinkage of the trees still to come (a forest keeps 1, DART recomputes its own every iteration)
    if (boost_ != kRF) shrinkage_rate_ = cfg_.learning_rate;
  }
}

// ---------------------------------------------------------------------------------------------------- objectives
void Booster::InitObjective() {
  const std::vector<float>& label = ds_->label;
  const double kEpsD = 1e-15f;
  if (cfg_.objective == "binary") {
    obj_ = kBinary;
    double cnt[2] = {0.0, 0.0};
    for (float l : label) cnt[l > 0 ? 1 : 0] += 1.0;
    net_->HostAllReduceSum(cnt, 2);    // label counts are summed over machines (a shard with one class still trains)
    const double neg = cnt[0], pos = cnt[1];
    need_train_ = !(pos == 0 || neg == 0);
    if (!need_train_) LogWarning("Contains only one class");
    label_weights_[0] = label_weights_[1] = 1.0;
    if (cfg_.is_unbalance && pos > 0 && neg > 0) {
      if (pos > neg) { label_weights_[1] = 1.0; label_weights_[0] = pos / neg; }
      else { label_weights_[1] = neg / pos; label_weights_[0] = 1.0; }
    }
    label_weights_[1] *= cfg_.scale_pos_weight;
  } else if (cfg_.objective == "regression") {
    obj_ = kL2;
  } else if (cfg_.objective == "lambdarank") {
    obj_ = kLambdarank;
    LGB_CHECK(!ds_->query_boundaries.empty(), "lambdarank needs query information (a group column)");
    std::vector<double> label_gain = cfg_.label_gain;
    if (label_gain.empty()) for (int i = 0; i < 31; ++i) label_gain.push_back(static_cast<double>((1LL << i) - 1));
    const int kMaxPosition = 10000;
    std::vector<double> discount(kMaxPosition);
    for (int i = 0; i < kMaxPosition; ++i) discount[i] = 1.0 / std::log2(2.0 + i);
    const auto& qb = ds_->query_boundaries;
    const int nq = static_cast<int>(qb.size()) - 1;
    std::vector<double> inv(nq);
    max_query_len_ = 0;
    for (float l : label) LGB_CHECK(l >= 0 && static_cast<int>(l) < static_cast<int>(label_gain.size()) && l == static_cast<int>(l),
                                    "lambdarank labels must be integers in [0, %d)", static_cast<int>(label_gain.size()));
    for (int q = 0; q < nq; ++q) {
      const int b = qb[q], cnt = qb[q + 1] - b;
      LGB_CHECK(cnt <= kMaxPosition, "A query has %d rows; the limit is %d", cnt, kMaxPosition);
      max_query_len_ = std::max(max_query_len_, cnt);
      // CalMaxDCGAtK
      std::vector<int> label_cnt(label_gain.size(), 0);
      for (int i = 0; i < cnt; ++i) ++label_cnt[static_cast<int>(label[b + i])];
      int top_label = static_cast<int>(label_gain.size()) - 1;
      int k = std::min(cfg_.lambdarank_truncation_level, cnt);
      double ret = 0.0;
      for (int j = 0; j < k; ++j) {
        while (top_label > 0 && label_cnt[top_label] <= 0) top_label -= 1;
        if (top_label < 0) break;
        ret += discount[j] * label_gain[top_label];
        label_cnt[top_label] -= 1;
      }
      if (ret > 0.0) ret = 1.0f / ret;
      inv[q] = ret;
    }
    // launch order of the gradient kernel: queries longer than short_query_len_ first (longest first), then the rest
    {
      std::vector<int> order(nq);
      for (int q = 0; q < nq; ++q) order[q] = q;
      std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return qb[a + 1] - qb[a] > qb[b + 1] - qb[b]; });
      n_long_queries_ = 0;
      while (n_long_queries_ < nq && qb[order[n_long_queries_] + 1] - qb[order[n_long_queries_]] > short_query_len_) ++n_long_queries_;
      d_query_order_.Alloc(std::max(nq, 1));
      LGB_CUDA(cudaMemcpy(d_query_order_.p, order.data(), sizeof(int) * nq, cudaMemcpyHostToDevice));
    }
    const size_t bins = 1024 * 1024;
    min_sigmoid_input_ = -50.0 / cfg_.sigmoid / 2;
    max_sigmoid_input_ = -min_sigmoid_input_;
    sigmoid_table_idx_factor_ = bins / (max_sigmoid_input_ - min_sigmoid_input_);
    std::vector<double> table(bins);
    for (size_t i = 0; i < bins; ++i) {
      const double sc = i / sigmoid_table_idx_factor_ + min_sigmoid_input_;
      table[i] = 1.0f / (1.0f + std::exp(sc * cfg_.sigmoid));
    }
    d_sigmoid_table_.Alloc(bins); LGB_CUDA(cudaMemcpy(d_sigmoid_table_.p, table.data(), sizeof(double) * bins, cudaMemcpyHostToDevice));
    d_inv_max_dcg_.Alloc(nq); LGB_CUDA(cudaMemcpy(d_inv_max_dcg_.p, inv.data(), sizeof(double) * nq, cudaMemcpyHostToDevice));
    d_label_gain_.Alloc(label_gain.size()); LGB_CUDA(cudaMemcpy(d_label_gain_.p, label_gain.data(), sizeof(double) * label_gain.size(), cudaMemcpyHostToDevice));
    d_discount_.Alloc(kMaxPosition); LGB_CUDA(cudaMemcpy(d_discount_.p, discount.data(), sizeof(double) * kMaxPosition, cudaMemcpyHostToDevice));
  } else if (cfg_.objective == "huber" || cfg_.objective == "fair" || cfg_.objective == "poisson" || cfg_.objective == "gamma" ||
             cfg_.objective == "tweedie") {
    const std::string& o = cfg_.objective;
    obj_ = o == "huber" ? kHuber : o == "fair" ? kFair : o == "poisson" ? kPoisson : o == "gamma" ? kGamma : kTweedie;
    if (LogLink()) {      // RegressionPoissonLoss::Init: the label domain of the log-link objectives
      double bad = 0.0, sum = 0.0;
      for (float l : label) { bad += (l < 0.0f || (obj_ == kGamma && l <= 0.0f)); sum += l; }
      double acc[2] = {bad, sum};
      net_->HostAllReduceSum(acc, 2);
      LGB_CHECK(acc[0] == 0.0, "objective=%s needs %s labels", o.c_str(), obj_ == kGamma ? "positive" : "non-negative");
      LGB_CHECK(acc[1] > 0.0, "objective=%s: the sum of the labels is zero", o.c_str());
    }
  } else if (cfg_.objective == "regression_l1" || cfg_.objective == "quantile" || cfg_.objective == "mape") {
    obj_ = cfg_.objective == "regression_l1" ? kL1 : cfg_.objective == "quantile" ? kQuantile : kMape;
    if (obj_ == kMape) {      // RegressionMAPELOSS::Init: label_weight = 1 / max(1, |label|) (* weight)
      label_weight_.resize(n_);
      for (int64_t i = 0; i < n_; ++i) {
        label_weight_[i] = 1.0f / std::max(1.0f, std::fabs(label[i]));
        if (has_weights_) label_weight_[i] = label_weight_[i] * ds_->weight[i];
      }
      d_label_weight_.Alloc(std::max<int64_t>(n_, 1));
      LGB_CUDA(cudaMemcpy(d_label_weight_.p, label_weight_.data(), sizeof(float) * n_, cudaMemcpyHostToDevice));
    }
  } else if (cfg_.objective == "multiclass") {
    // MulticlassSoftmax::Init (upstream multiclass_objective.hpp): label shares over all machines, classes that never / always occur
    obj_ = kMulticlass;
    K_ = cfg_.num_class;
    LGB_CHECK(K_ >= 2 && K_ <= kMaxClasses, "objective=multiclass needs 2 <= num_class <= %d (got %d)", kMaxClasses, K_);
    std::vector<double> acc(K_ + 1, 0.0);      // [K_] = sum of weights
    const std::vector<float>& wv = ds_->weight;
    for (int64_t i = 0; i < n_; ++i) {
      const int l = static_cast<int>(label[i]);
      LGB_CHECK(l >= 0 && l < K_ && label[i] == static_cast<float>(l), "multiclass labels must be integers in [0, %d)", K_);
      const double w = has_weights_ ? wv[i] : 1.0;
      acc[l] += w; acc[K_] += w;
    }
    net_->HostAllReduceSum(acc.data(), K_ + 1);
    class_init_probs_.assign(K_, 0.0);
    class_need_train_.assign(K_, 1);
    for (int k = 0; k < K_; ++k) {
      class_init_probs_[k] = acc[k] / acc[K_];
      if (std::fabs(class_init_probs_[k]) <= kEpsD || std::fabs(class_init_probs_[k]) >= 1.0 - kEpsD) class_need_train_[k] = 0;
    }
  } else if (cfg_.objective == "cross_entropy") {
    obj_ = kXentropy;      // CrossEntropy (upstream xentropy_objective.hpp): labels are probabilities in [0, 1]
    static_assert(kXentropy == 15, "k_grad_point's kind for cross_entropy");
    for (float l : label) LGB_CHECK(l >= 0.0f && l <= 1.0f, "objective=cross_entropy needs labels in [0, 1]");
  } else if (cfg_.objective == "rank_xendcg") {
    // RankXENDCG::Init (upstream rank_objective.hpp): one random stream per query of this machine, seeded objective_seed + query
    obj_ = kXendcg;
    LGB_CHECK(!ds_->query_boundaries.empty(), "rank_xendcg needs query information (a group column)");
    const int nq = static_cast<int>(ds_->query_boundaries.size()) - 1;
    std::vector<unsigned int> st(std::max(nq, 1));
    max_query_len_ = 1;
    for (int q = 0; q < nq; ++q) {
      st[q] = static_cast<unsigned int>(cfg_.objective_seed + q);
      max_query_len_ = std::max(max_query_len_, ds_->query_boundaries[q + 1] - ds_->query_boundaries[q]);
    }
    for (float l : label) LGB_CHECK(l >= 0 && l < 31 && l == static_cast<int>(l), "rank_xendcg labels must be integers in [0, 31)");
    d_xe_rand_.Alloc(st.size());
    LGB_CUDA(cudaMemcpy(d_xe_rand_.p, st.data(), sizeof(unsigned int) * st.size(), cudaMemcpyHostToDevice));
  } else if (cfg_.objective == "multiclassova") {
    // MulticlassOVA::Init: one BinaryLogloss per class whose positives are the rows of that class (label counts, is_unbalance /
    // scale_pos_weight weights and the sums behind BoostFromScore, all over all machines)
    obj_ = kMulticlassOva;
    K_ = cfg_.num_class;
    LGB_CHECK(K_ >= 2 && K_ <= kMaxClasses, "objective=multiclassova needs 2 <= num_class <= %d (got %d)", kMaxClasses, K_);
    std::vector<double> acc(2 * K_ + 2, 0.0);      // [k]: rows of class k, [K + k]: their weight, [2K]: rows, [2K + 1]: weight
    const std::vector<float>& wv = ds_->weight;
    for (int64_t i = 0; i < n_; ++i) {
      const int l = static_cast<int>(label[i]);
      LGB_CHECK(l >= 0 && l < K_ && label[i] == static_cast<float>(l), "multiclassova labels must be integers in [0, %d)", K_);
      const double w = has_weights_ ? wv[i] : 1.0;
      acc[l] += 1.0; acc[K_ + l] += w; acc[2 * K_] += 1.0; acc[2 * K_ + 1] += w;
    }
    net_->HostAllReduceSum(acc.data(), 2 * K_ + 2);
    class_need_train_.assign(K_, 1);
    ova_label_weights_.assign(2 * K_, 1.0);
    ova_suml_.assign(K_, 0.0);
    ova_sumw_ = acc[2 * K_ + 1];
    for (int k = 0; k < K_; ++k) {
      const double pos = acc[k], neg = acc[2 * K_] - acc[k];
      if (pos == 0 || neg == 0) class_need_train_[k] = 0;
      if (cfg_.is_unbalance && pos > 0 && neg > 0) {
        if (pos > neg) ova_label_weights_[2 * k] = pos / neg;
        else ova_label_weights_[2 * k + 1] = neg / pos;
      }
      ova_label_weights_[2 * k + 1] *= cfg_.scale_pos_weight;
      ova_suml_[k] = acc[K_ + k];
    }
  } else if (cfg_.objective == "custom" || cfg_.objective == "none" || cfg_.objective == "null" || cfg_.objective == "na") {
    obj_ = kCustom;      // no objective function: LGBM_BoosterUpdateOneIterCustom supplies the gradients, scores stay raw
    K_ = std::max(1, cfg_.num_class);      // num_tree_per_iteration = num_class also without an objective (GBDT::Init): the caller's arrays
                                           // then hold one value per row and class, class after class (FObjTrait with a multiclass label)
    LGB_CHECK(K_ <= kMaxClasses, "num_class=%d is above %d", K_, kMaxClasses);
    if (K_ > 1) class_need_train_.assign(K_, 1);
  } else {
    throw std::runtime_error("objective=" + cfg_.objective + " is not supported by this library yet (binary, cross_entropy, multiclass, multiclassova, rank_xendcg, regression, regression_l1, quantile, mape, huber, fair, poisson, gamma, tweedie, lambdarank, custom)");
  }
}

double Booster::BoostFromScoreLocalSums(double* suml, double* sumw) const {
  const std::vector<float>& label = ds_->label;
  const std::vector<float>& w = ds_->weight;
  double sl = 0.0, sw = 0.0;
  if (obj_ == kBinary) {
    if (has_weights_) { for (int64_t i = 0; i < n_; ++i) { sl += (label[i] > 0) * w[i]; sw += w[i]; } }
    else { sw = static_cast<double>(n_); for (int64_t i = 0; i < n_; ++i) sl += (label[i] > 0); }
  } else {
    if (has_weights_) { for (int64_t i = 0; i < n_; ++i) { sl += label[i] * w[i]; sw += w[i]; } }
    else { sw = static_cast<double>(n_); for (int64_t i = 0; i < n_; ++i) sl += label[i]; }
  }
  *suml = sl; *sumw = sw;
  return 0.0;
}

void Booster::ComputeGradients(const double* score_override) {
  const int n = static_cast<int>(n_);
  const double* score = score_override ? score_override : d_score_.p;
  const int grid = std::min(CeilDiv(n, 256), kNumSms * 8);
  const float* w = has_weights_ ? ds_->d_weight.p : nullptr;
  if (obj_ == kMulticlass) {
    k_grad_multiclass<<<grid, 256, 0, stream_>>>(score, ds_->d_label.p, w, d_grad_[0].p, d_hess_[0].p, n, K_, static_cast<double>(K_) / (K_ - 1.0f));
  } else if (obj_ == kXendcg) {
    XendcgArgs a;
    a.score = score; a.label = ds_->d_label.p; a.weight = w; a.qb = ds_->d_query_boundaries.p; a.rand_state = d_xe_rand_.p;
    a.grad = d_grad_[grad_cur_].p; a.hess = d_hess_[grad_cur_].p; a.max_cnt = max_query_len_;
    const size_t smem = static_cast<size_t>(max_query_len_) * 2 * sizeof(double);
    LGB_CHECK(smem <= 200 * 1024, "rank_xendcg: a query with %d documents does not fit the kernel's shared memory", max_query_len_);
    static size_t smem_set = 0;
    if (smem > smem_set) { smem_set = smem; SetSmem(k_grad_xendcg, smem_set); }
    const int nq = static_cast<int>(ds_->query_boundaries.size()) - 1;
    if (nq > 0) k_grad_xendcg<<<nq, 256, smem, stream_>>>(a);
  } else if (obj_ == kMulticlassOva) {
    for (int k = 0; k < K_; ++k)
      k_grad_binary<<<grid, 256, 0, stream_>>>(score + static_cast<size_t>(k) * n_, ds_->d_label.p, w, d_grad_[0].p + static_cast<size_t>(k) * n_,
                                               d_hess_[0].p + static_cast<size_t>(k) * n_, n, cfg_.sigmoid, ova_label_weights_[2 * k],
                                               ova_label_weights_[2 * k + 1], k);
    launches_ += K_ - 1;
  } else if (obj_ == kBinary) {
    k_grad_binary<<<grid, 256, 0, stream_>>>(score, ds_->d_label.p, w, d_grad_[grad_cur_].p, d_hess_[grad_cur_].p, n, cfg_.sigmoid, label_weights_[0], label_weights_[1]);
  } else if (obj_ == kL2) {
    k_grad_l2<<<grid, 256, 0, stream_>>>(score, ds_->d_label.p, w, d_grad_[grad_cur_].p, d_hess_[grad_cur_].p, n);
  } else if (obj_ >= kHuber) {      // (kXentropy included: kind 15 of k_grad_point)
    const double p0 = (obj_ == kHuber || obj_ == kQuantile) ? cfg_.alpha : obj_ == kFair ? cfg_.fair_c
                      : obj_ == kPoisson ? SoftExp(cfg_.poisson_max_delta_step) : cfg_.tweedie_variance_power;
    k_grad_point<<<grid, 256, 0, stream_>>>(static_cast<int>(obj_), score, ds_->d_label.p, w, d_grad_[grad_cur_].p, d_hess_[grad_cur_].p, n, p0,
                                            d_label_weight_.p);
  } else {
    LambdarankArgs a;
    a.score = score; a.label = ds_->d_label.p; a.weight = w; a.qb = ds_->d_query_boundaries.p;
    a.inv_max_dcg = d_inv_max_dcg_.p; a.label_gain = d_label_gain_.p; a.discount = d_discount_.p; a.sigmoid_table = d_sigmoid_table_.p;
    a.grad = d_grad_[grad_cur_].p; a.hess = d_hess_[grad_cur_].p; a.sigmoid = cfg_.sigmoid; a.min_sigmoid_input = min_sigmoid_input_;
    a.max_sigmoid_input = max_sigmoid_input_; a.idx_factor = sigmoid_table_idx_factor_;
    a.truncation_level = cfg_.lambdarank_truncation_level; a.norm = cfg_.lambdarank_norm; a.table_size = 1024 * 1024;
    auto smem_for = [](int max_cnt) { return static_cast<size_t>(max_cnt) * (2 * sizeof(double) + 5 * sizeof(float) + sizeof(int)) + 16; };
    // the attribute is the LIMIT of what a launch may ask for: it must cover the short-query launch too (it asks for short_query_len_
    // documents even when the longest query is shorter — a limit set from the longest query alone refused that launch)
    static size_t smem_set = 0;
    const size_t smem_need = std::max(smem_for(max_query_len_), smem_for(short_query_len_));
    if (smem_need > smem_set) { smem_set = smem_need; SetSmem(k_grad_lambdarank, smem_set); }
    // long queries first (they are the tail), then the short ones with a small shared-memory footprint (more CTAs per SM)
    if (n_long_queries_ > 0) {
      a.max_cnt = max_query_len_; a.qlist = d_query_order_.p;
      k_grad_lambdarank<<<n_long_queries_, 256, smem_for(max_query_len_), stream_>>>(a);
    }
    const int nq = static_cast<int>(ds_->query_boundaries.size()) - 1;
    if (nq > n_long_queries_) {
      a.max_cnt = short_query_len_; a.qlist = d_query_order_.p + n_long_queries_;
      k_grad_lambdarank<<<nq - n_long_queries_, 256, smem_for(short_query_len_), stream_>>>(a);
    }
  }
  LGB_CUDA(cudaGetLastError());
  ++launches_;
}

void Booster::PrepareGradients(bool absmax_ready) {
  const int n = static_cast<int>(n_);
  const int grid = std::min(CeilDiv(n, 256), kNumSms * 8);
  if (!absmax_ready) {
    LGB_CUDA(cudaMemsetAsync(&d_ctrl_->max_abs_g_bits, 0, 2 * sizeof(unsigned int), stream_));
    k_absmax<<<grid, 256, 0, stream_>>>(d_grad_[grad_cur_].p + cls_ * n_, d_hess_[grad_cur_].p + cls_ * n_, n, d_ctrl_);
  }
  if (peers_.world > 0) { k_peer_max<<<1, 32, 0, stream_>>>(d_ctrl_, peers_, ++max_seq_); ++launches_; }
  else if (nccl_exchange_) net_->AllReduceMaxU32(&d_ctrl_->max_abs_g_bits, 2, stream_);
  k_scales<<<1, 1, 0, stream_>>>(d_ctrl_, lp_);
  // the stochastic-rounding stream is keyed by the tree (iteration * K + class) and the global row
  k_quantize<<<grid, 256, 0, stream_>>>(d_grad_[grad_cur_].p + cls_ * n_, d_hess_[grad_cur_].p + cls_ * n_, n, static_cast<const BatchCtrl*>(d_ctrl_), lp_,
                                        static_cast<unsigned int>(iter_ * K_ + cls_),
                                        static_cast<unsigned int>(row_offset_), d_gh16_.p, d_ghfix_.p,
                                        bagging_ ? d_bag_bits_.p : nullptr);
  LGB_CUDA(cudaGetLastError());
  launches_ += 3;
}

// ---------------------------------------------------------------------------------------------------- tree learner
void Booster::LaunchHist(bool root) {
  const BinLayout& lay = ds_->layout;
  const int R = lay.rows_per_step;
  const bool flt = !lp_.quantized;
  const int ft = flt ? lay.tile_features_f : lay.tile_features_q;
  const int tiles = CeilDiv(F_, ft);
  const int row_ctas = std::max(1, kNumSms / tiles);
  dim3 grid(row_ctas, tiles);
  int rows_per_flush;
  if (flt) {
    rows_per_flush = kFixRowsPerCopy * R;
  } else {
    const int qb = cfg_.num_grad_quant_bins;
    const int per_copy = std::min(32767 / (qb / 2 + 1), 65535 / (qb + 1));
    rows_per_flush = std::max(32, (per_copy * R) / 32 * 32);
  }
  // non-root launches find their work areas already cleared by the partition kernel; the work areas alternate between two
  // buffers from one histogram launch to the next (what the peer exchange needs: batch_tree.cuh)
  PeerTable pt = peers_;
  ++exchange_seq_;
  pt.seq = exchange_seq_;
  long long* work = d_hist_work_.p + (exchange_seq_ & 1u) * work_stride_;
  if (root) LGB_CUDA(cudaMemsetAsync(work, 0, sizeof(long long) * work_elems_, stream_));
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  if (time_hist_) {
    if (hist_events_total_ == hist_events_.size()) {
      cudaEvent_t a, b;
      LGB_CUDA(cudaEventCreate(&a)); LGB_CUDA(cudaEventCreate(&b));
      hist_events_.emplace_back(a, b);
    }
    e0 = hist_events_[hist_events_total_].first; e1 = hist_events_[hist_events_total_].second;
    ++hist_events_total_;
    LGB_CUDA(cudaEventRecord(e0, stream_));
  }
#define LAUNCH_HIST_ARGS ds_->d_rm.p, static_cast<unsigned>(lay.rm_stride), d_gh16_.p, d_ghfix_.p, d_idx_[0].p, d_idx_[1].p, \
                         d_nodes_, d_ctrl_, work, work_elems_, F_, rows_per_flush, row_ctas, lp_, pt
#define LAUNCH_HIST(RR, BPL, FLT)                                                                                     \
  do {                                                                                                                \
    if (root) LaunchPdl(k_bt_hist<RR, BPL, FLT, true>, grid, dim3(kHistThreads), kHistSmem, stream_, LAUNCH_HIST_ARGS);  \
    else LaunchPdl(k_bt_hist<RR, BPL, FLT, false>, grid, dim3(kHistThreads), kHistSmem, stream_, LAUNCH_HIST_ARGS);      \
  } while (0)
  if (!flt) {
    if (R == 4) LAUNCH_HIST(4, 4, false); else if (R == 2) LAUNCH_HIST(2, 4, false); else LAUNCH_HIST(1, 4, false);
  } else {
    if (R == 4) LAUNCH_HIST(4, 2, true); else if (R == 2) LAUNCH_HIST(2, 2, true); else LAUNCH_HIST(1, 2, true);
  }
#undef LAUNCH_HIST
#undef LAUNCH_HIST_ARGS
  LGB_CUDA(cudaGetLastError());
  if (time_hist_) LGB_CUDA(cudaEventRecord(e1, stream_));
  ++launches_;
  ++hist_kernel_launches_;
  // data-parallel learner: every rank needs the global histograms of the batch (integer sums: order-independent)
  if (nccl_exchange_)
    net_->AllReduceSumInt64(work, static_cast<size_t>(work_elems_) * (root ? 1 : max_batch_), stream_);
}

void Booster::LaunchFind(bool root) {
  const int funits = peers_.sliced ? CeilDiv(F_, peers_.world) : F_;      // sliced: the largest feature range of any rank
  PeerTable pt = peers_;
  pt.seq = exchange_seq_;
  pt.parity_off = (exchange_seq_ & 1u) * work_stride_;
  long long* work = d_hist_work_.p + (exchange_seq_ & 1u) * work_stride_;
  // narrow data: a 256-thread CTA per (feature, child, batch item) — a few hundred CTAs, lowest latency; wide data: a warp per unit
  if (!find_warp_) {
    dim3 grid(funits, root ? 1 : 2, root ? 1 : max_batch_);
    // (the second template flag compiles the categorical / wide-feature / constrained search in or out: batch_tree.cuh)
    const bool ext = ds_->has_categorical || ds_->has_wide || lp_.monotone != nullptr || lp_.node_mask != nullptr;
#define LGB_FIND(Q, E) LaunchPdl(k_bt_find<Q, E>, grid, dim3(256), 0, stream_, work, work_elems_, d_hist_pool_.p, ds_->d_fmeta.p, d_nodes_, d_ctrl_, \
                                 d_feat_best_.p, d_splittable_.p, lp_, root ? 1 : 0, pt)
    if (lp_.quantized) { if (ext) LGB_FIND(true, true); else LGB_FIND(true, false); }
    else { if (ext) LGB_FIND(false, true); else LGB_FIND(false, false); }
#undef LGB_FIND
  } else {
    dim3 grid(CeilDiv(funits, kFindWarps), root ? 1 : 2, root ? 1 : max_batch_);
    if (lp_.quantized)
      LaunchPdl(k_bt_find_warp<true>, grid, dim3(kFindWarps * 32), 0, stream_, work, work_elems_, d_hist_pool_.p, ds_->d_fmeta.p, d_nodes_, d_ctrl_,
                d_feat_best_.p, d_splittable_.p, lp_, root ? 1 : 0, pt);
    else
      LaunchPdl(k_bt_find_warp<false>, grid, dim3(kFindWarps * 32), 0, stream_, work, work_elems_, d_hist_pool_.p, ds_->d_fmeta.p, d_nodes_, d_ctrl_,
                d_feat_best_.p, d_splittable_.p, lp_, root ? 1 : 0, pt);
  }
  ++launches_;
}

// Replay + next batch.  (Running it inside the split-search kernel's last sub-block was measured: no faster — the replay's own
// ~10 us stay on the critical path and the search kernel loses registers.)
void Booster::LaunchSelect(bool after_root) {
  ProfBegin("select");
  PeerTable pt = peers_;
  pt.seq = exchange_seq_;
  LaunchPdl(k_bt_select, dim3(1), dim3(256), SelectSmemBytes(cfg_.num_leaves), stream_, d_nodes_, d_avail_.p, d_ctrl_, tree_dev_, ds_->d_fmeta.p,
            lp_, max_nodes_, spec_budget_, max_batch_, static_cast<const SplitDev*>(d_feat_best_.p), after_root ? 1 : 0, pt);
  ProfEnd();
  ++launches_;
}

// One round = expand every leaf of the current batch (partition, histogram of the smaller child, split search for both
// children), then replay the sequential order and choose the next batch.  Every kernel reads its work from device memory
// and returns at once when the tree is finished, so rounds are enqueued without waiting for the GPU.
void Booster::EnqueueRounds(int rounds) {
  for (int r = 0; r < rounds; ++r) {
    ProfBegin("partition");
    // the partition kernel clears the work areas the round's histograms go into
    long long* next_work = d_hist_work_.p + ((exchange_seq_ + 1u) & 1u) * work_stride_;
    if (idx_es_ == 2)
      LaunchPdl(k_bt_partition<2>, dim3(part_blocks_), dim3(kPartThreads), 0, stream_, ds_->d_cm.p, static_cast<long long>(ds_->n_pad),
                d_idx_[0].p, d_idx_[1].p, d_idx_[0].p, d_idx_[1].p, d_gh16_.p, d_ghfix_.p, d_ctrl_, next_work, work_elems_);
    else
      LaunchPdl(k_bt_partition<4>, dim3(part_blocks_), dim3(kPartThreads), 0, stream_, ds_->d_cm.p, static_cast<long long>(ds_->n_pad),
                d_idx_[0].p, d_idx_[1].p, d_idx_[0].p, d_idx_[1].p, d_gh16_.p, d_ghfix_.p, d_ctrl_, next_work, work_elems_);
    ProfEnd();
    ProfBegin("hist"); LaunchHist(false); ProfEnd();
    launches_ += 1;
    ProfBegin("find"); LaunchFind(false); ProfEnd();
    LaunchSelect(false);
  }
}

// ColSampler::ResetByTree (upstream src/treelearner/col_sampler.hpp): every tree draws max(round(F * feature_fraction), min(2, F))
// of the used features from one LCG stream; every rank draws the same (same seed, same F).  The split search treats the others
// as if they had no valid threshold.
// BaggingSampleStrategy::Bagging: a new bag at every iteration that is a multiple of bagging_freq (each rank over its own rows).
// The root's row count is the number of rows in the bags of all ranks.
void Booster::Bagging() {
  if (boost_ == kGOSS) { GossSample(); return; }
  if (!bagging_ || iter_ % cfg_.bagging_freq != 0) return;
  const int n = static_cast<int>(n_);
  LGB_CUDA(cudaMemsetAsync(d_bag_count_.p, 0, sizeof(int), stream_));
  if (balanced_bagging_)
    k_bag_sample<<<CeilDiv(CeilDiv(n, 1024), 64), 64, 0, stream_>>>(d_bag_rand_.p, n, cfg_.neg_bagging_fraction, d_bag_bits_.p, d_bag_count_.p,
                                                                    ds_->d_label.p, cfg_.pos_bagging_fraction);
  else
    k_bag_sample<<<CeilDiv(CeilDiv(n, 1024), 64), 64, 0, stream_>>>(d_bag_rand_.p, n, cfg_.bagging_fraction, d_bag_bits_.p, d_bag_count_.p);
  LGB_CUDA(cudaGetLastError());
  ++launches_;
  int local = 0;
  LGB_CUDA(cudaMemcpyAsync(&local, d_bag_count_.p, sizeof(int), cudaMemcpyDeviceToHost, stream_));
  LGB_CUDA(cudaStreamSynchronize(stream_));
  std::vector<long long> all;
  net_->HostAllGatherInt64(local, &all);
  bag_cnt_global_ = 0;
  for (long long c : all) bag_cnt_global_ += c;
}

// GOSS (upstream src/boosting/goss.hpp): from iteration 1 / learning_rate on, every 1024-row block keeps its top_rate rows by
// |g * h| and a random other_rate of the rest with scaled-up gradients (k_goss_sample); before that every row is in the bag.
void Booster::GossSample() {
  const int n = static_cast<int>(n_);
  const int nblk = CeilDiv(n, 1024);
  int local = n;
  if (iter_ < static_cast<int>(1.0f / cfg_.learning_rate)) {
    k_bag_all<<<CeilDiv(nblk * 32, 256), 256, 0, stream_>>>(d_bag_bits_.p, n);
  } else {
    LGB_CUDA(cudaMemsetAsync(d_bag_count_.p, 0, sizeof(int), stream_));
    k_goss_sample<<<nblk, 256, 0, stream_>>>(d_grad_[grad_cur_].p, d_hess_[grad_cur_].p, n, K_, cfg_.top_rate, cfg_.other_rate, d_bag_rand_.p,
                                             d_bag_bits_.p, d_bag_count_.p);
    LGB_CUDA(cudaMemcpyAsync(&local, d_bag_count_.p, sizeof(int), cudaMemcpyDeviceToHost, stream_));
    LGB_CUDA(cudaStreamSynchronize(stream_));
  }
  LGB_CUDA(cudaGetLastError());
  ++launches_;
  std::vector<long long> all;
  net_->HostAllGatherInt64(local, &all);
  bag_cnt_global_ = 0;
  for (long long c : all) bag_cnt_global_ += c;
}

// ScoreUpdater::AddScore(tree, class) for an earlier tree: its arrays go to the second device tree and every row walks it.
void Booster::AddHostTreeToScores(const HostTree& t, int cls, bool train, bool valid) {
  const int L = cfg_.num_leaves;
  const int nl = t.num_leaves;
  LGB_CHECK(nl <= L, "AddHostTreeToScores: the tree has more leaves than num_leaves");
  LGB_CHECK(nl <= 1 || static_cast<int>(t.split_feature_inner.size()) == nl - 1, "this tree was not trained by this booster (no bin-level splits)");
  unsigned char* b = h_tree_blob2_.data();
  const size_t dbl = 4 * sizeof(double) * L, ints = 8 * sizeof(int) * L, floats = RoundUp<size_t>(sizeof(float) * L, 8);
  double* dv = reinterpret_cast<double*>(b);
  int* ip = reinterpret_cast<int*>(b + dbl);
  signed char* cp = reinterpret_cast<signed char*>(b + dbl + ints + floats);
  for (int i = 0; i < nl; ++i) dv[i] = t.leaf_value[i];
  for (int i = 0; i < nl - 1; ++i) {
    ip[i] = t.left_child[i]; ip[L + i] = t.right_child[i]; ip[2 * L + i] = t.split_feature_inner[i]; ip[3 * L + i] = t.threshold_in_bin[i];
    cp[i] = t.decision_type[i];
  }
  if (ds_->has_categorical) {
    unsigned int* cb = reinterpret_cast<unsigned int*>(b + tree_cat_off_);
    LGB_CHECK(t.num_cat == 0 || static_cast<int>(t.cat_bits_inner.size()) == 8 * (nl - 1), "this tree was not trained by this booster (no bin-level category sets)");
    for (int i = 0; i < 8 * (nl - 1); ++i) cb[i] = t.num_cat > 0 ? t.cat_bits_inner[i] : 0u;
  }
  LGB_CUDA(cudaMemcpyAsync(d_tree_blob2_.p, b, tree_blob_bytes_, cudaMemcpyHostToDevice, stream_));
  LGB_CUDA(cudaStreamSynchronize(stream_));      // the host blob is reused by the next tree
  const ScoreMix plain{1.0, 1.0, 0.0, 0};
  auto run = [&](const Dataset* d, double* score) {
    const int n = static_cast<int>(d->num_data);
    if (n == 0) return;
    const BinLayout& lay = d->layout;
    const int grid = std::min(CeilDiv(n, kApplyThreads), kNumSms * 8);
#define APPLY_P(WPR, CAT) k_bt_apply_tree<0, WPR, CAT><<<grid, kApplyThreads, 0, stream_>>>(d->d_rm.p, static_cast<unsigned>(lay.rm_stride), score, \
        nullptr, nullptr, nullptr, nullptr, tree_dev2_, ds_->d_fmeta.p, static_cast<BatchCtrl*>(nullptr), 1.0, n, cfg_.sigmoid, 1.0, 1.0, plain, nl)
    if (ds_->has_categorical) { if (lay.rm_stride == 32) APPLY_P(8, true); else if (lay.rm_stride == 64) APPLY_P(16, true); else APPLY_P(0, true); }
    else if (lay.rm_stride == 32) APPLY_P(8, false); else if (lay.rm_stride == 64) APPLY_P(16, false); else APPLY_P(0, false);
#undef APPLY_P
    ++launches_;
  };
  if (train) run(ds_, d_score_.p + static_cast<size_t>(cls) * n_);
  if (valid) for (ValidSet& v : valids_) run(v.ds, v.score.p + static_cast<size_t>(cls) * v.ds->num_data);
  LGB_CUDA(cudaGetLastError());
}

// DART::DroppingTrees (upstream src/boosting/dart.hpp): draws the iterations to drop, takes their trees out of the training scores
// and sets this iteration's shrinkage.
void Booster::DroppingTrees() {
  drop_index_.clear();
  const bool is_skip = drop_rand_.NextFloat() < cfg_.skip_drop;
  if (!is_skip) {
    double drop_rate = cfg_.drop_rate;
    if (!cfg_.uniform_drop) {
      const double inv_average_weight = static_cast<double>(tree_weight_.size()) / sum_weight_;
      if (cfg_.max_drop > 0) drop_rate = std::min(drop_rate, cfg_.max_drop * inv_average_weight / sum_weight_);
      for (int i = 0; i < iter_; ++i) {
        if (drop_rand_.NextFloat() < drop_rate * tree_weight_[i] * inv_average_weight) {
          drop_index_.push_back(num_init_iteration_ + i);
          if (drop_index_.size() >= static_cast<size_t>(cfg_.max_drop)) break;
        }
      }
    } else {
      if (cfg_.max_drop > 0) drop_rate = std::min(drop_rate, cfg_.max_drop / static_cast<double>(iter_));
      for (int i = 0; i < iter_; ++i) {
        if (drop_rand_.NextFloat() < drop_rate) {
          drop_index_.push_back(num_init_iteration_ + i);
          if (drop_index_.size() >= static_cast<size_t>(cfg_.max_drop)) break;
        }
      }
    }
  }
  for (int i : drop_index_) {
    for (int k = 0; k < K_; ++k) {
      HostTree& t = models_[static_cast<size_t>(i) * K_ + k];
      t.Shrinkage(-1.0);
      AddHostTreeToScores(t, k, true, false);
    }
  }
  if (!cfg_.xgboost_dart_mode) shrinkage_rate_ = cfg_.learning_rate / (1.0f + static_cast<double>(drop_index_.size()));
  else if (drop_index_.empty()) shrinkage_rate_ = cfg_.learning_rate;
  else shrinkage_rate_ = cfg_.learning_rate / (cfg_.learning_rate + static_cast<double>(drop_index_.size()));
}

// DART::Normalize: the dropped trees come back scaled by k / (k + 1) (xgboost mode: k / (k + learning_rate)); the validation
// scores, which kept them, lose the difference.
void Booster::NormalizeDart() {
  const double k = static_cast<double>(drop_index_.size());
  for (int i : drop_index_) {
    for (int c = 0; c < K_; ++c) {
      HostTree& t = models_[static_cast<size_t>(i) * K_ + c];
      if (!cfg_.xgboost_dart_mode) t.Shrinkage(1.0f / (k + 1.0f)); else t.Shrinkage(shrinkage_rate_);
      AddHostTreeToScores(t, c, false, true);
      if (!cfg_.xgboost_dart_mode) t.Shrinkage(-k); else t.Shrinkage(-k / cfg_.learning_rate);
      AddHostTreeToScores(t, c, true, false);
    }
    if (!cfg_.uniform_drop) {
      const int w = i - num_init_iteration_;
      if (!cfg_.xgboost_dart_mode) {
        sum_weight_ -= tree_weight_[w] * (1.0f / (k + 1.0f));
        tree_weight_[w] *= (k / (k + 1.0f));
      } else {
        sum_weight_ -= tree_weight_[w] * (1.0f / (k + cfg_.learning_rate));
        tree_weight_[w] *= (k / (k + cfg_.learning_rate));
      }
    }
  }
  ++model_version_;
}

void Booster::SampleFeaturesForTree() {
  // ColSampler::ResetByTree draws over the used FEATURES; a wide feature's columns share its draw
  const int NF = ds_->num_logical_features();
  const int min_cnt = std::min(2, NF);
  const int cnt = std::max(static_cast<int>(NF * cfg_.feature_fraction + 0.5), min_cnt);
  std::vector<unsigned char>& mask = tree_mask_host_;
  mask.assign(F_, 0);
  for (int lf : feature_rand_.Sample(NF, cnt))
    for (int c = 0; c < ds_->logical_num_cols[lf]; ++c) mask[ds_->logical_first_col[lf] + c] = 1;
  // pageable source: the runtime stages it before the call returns, so the vector may go out of scope
  LGB_CUDA(cudaMemcpyAsync(d_feature_mask_.p, mask.data(), F_, cudaMemcpyHostToDevice, stream_));
}

// ColSampler::GetByNode for every search this tree can make (the root, then the smaller and the larger child of every split but
// the last), drawn from the tree's own column sample when there is one, on the column sampler's stream.  How many of them the tree
// really used is known when it is finished (FetchTree); the stream is then rewound and advanced by exactly that many draws.
void Booster::SampleNodeMasksForTree() {
  std::vector<int> pool;      // logical features
  for (int lf = 0; lf < ds_->num_logical_features(); ++lf)
    if (lp_.feature_mask == nullptr || tree_mask_host_[ds_->logical_first_col[lf]]) pool.push_back(lf);
  node_pool_size_ = static_cast<int>(pool.size());
  node_sample_cnt_ = std::max(static_cast<int>(node_pool_size_ * cfg_.feature_fraction_bynode + 0.5), std::min(2, node_pool_size_));
  node_rand_saved_ = feature_rand_;
  const int draws = 1 + 2 * (cfg_.num_leaves - 1);
  std::vector<unsigned char> masks(static_cast<size_t>(draws) * F_, 0);
  for (int d = 0; d < draws; ++d)
    for (int i : feature_rand_.Sample(node_pool_size_, node_sample_cnt_)) {
      const int lf = pool[i];
      for (int c = 0; c < ds_->logical_num_cols[lf]; ++c) masks[static_cast<size_t>(d) * F_ + ds_->logical_first_col[lf] + c] = 1;
    }
  LGB_CUDA(cudaMemcpyAsync(d_node_mask_.p, masks.data(), masks.size(), cudaMemcpyHostToDevice, stream_));
  LGB_CUDA(cudaStreamSynchronize(stream_));
}

void Booster::TrainTree(HostTree* out) {
  const int L = cfg_.num_leaves;
  const int n = static_cast<int>(n_);
  if (trace_cap_ > 0) {      // trace three trees starting at iteration LGB_B200_TRACE_FROM
    long long* p = (iter_ >= trace_from_ && iter_ < trace_from_ + 3) ? d_trace_.p : nullptr;
    LGB_CUDA(cudaMemcpyAsync(&d_ctrl_->trace, &p, sizeof(p), cudaMemcpyHostToDevice, stream_));
    const int on = p != nullptr ? 1 : 0;
    LGB_CUDA(cudaMemcpyToSymbolAsync(c_trace_on, &on, sizeof(int), 0, cudaMemcpyHostToDevice, stream_));
    LGB_CUDA(cudaStreamSynchronize(stream_));
  }
  if (lp_.feature_mask != nullptr) SampleFeaturesForTree();
  if (lp_.node_mask != nullptr) SampleNodeMasksForTree();
  k_bt_init<<<CeilDiv(L, 256), 256, 0, stream_>>>(d_nodes_, d_avail_.p, d_ctrl_, tree_dev_, L, n,
                                                  static_cast<int>(bagging_ ? bag_cnt_global_ : n_global_));
  LGB_CUDA(cudaMemsetAsync(d_splittable_.p, 1, d_splittable_.bytes(), stream_));
  ++launches_;
  ProfBegin("hist_root"); LaunchHist(true); ProfEnd();
  ProfBegin("find"); LaunchFind(true); ProfEnd();
  LaunchSelect(true);
  // The number of rounds a tree needs is only known on the device.  Rounds enqueued beyond it return at once but still cost a
  // launch each (~2.5 us per kernel, 4 kernels), a tree that is not finished costs a host round trip (~20 us): enqueue what the
  // previous tree used + round_slack_, then look, and top up two rounds at a time.
  int rounds = last_rounds_ > 0 ? std::min(L - 1, last_rounds_ + round_slack_) : std::min(L - 1, (L - 1) / 2 + 2);
  rounds = std::max(rounds, 1);
  int total = 0;
  const bool renew = RenewsLeaves();
  while (true) {
    EnqueueRounds(rounds);
    LGB_CUDA(cudaEventRecord(ev_rounds_done_, stream_));
    // (l1 / quantile / mape: the leaf values are re-fit before the tree may enter the scores)
    if (!renew) { ProfBegin("score_update"); UpdateScores(); ProfEnd(); }      // does nothing on the device unless the tree is finished
    total += rounds;
    LGB_CUDA(cudaGetLastError());
    if (FetchTree(out)) {
      if (lp_.node_mask != nullptr) {
        // upstream never searches the children of the split that fills the tree; here they were searched before that split was
        // committed: their two samples do not count
        const bool full = out->num_leaves >= L;
        const int used = last_node_draws_ - ((full && last_searched_) ? 2 : 0);
        feature_rand_ = node_rand_saved_;
        for (int d = 0; d < used; ++d) feature_rand_.Sample(node_pool_size_, node_sample_cnt_);
      }
      if (renew) { RenewTreeOutput(out); ProfBegin("score_update"); UpdateScores(); ProfEnd(); }
      break;
    }
    ++refetches_;
    LGB_CHECK(total < 2 * L, "tree growth did not finish");   // cannot happen: every round commits at least one split
    rounds = std::min(L - 1, 2);
  }
  rounds_enqueued_ += total;
}

void Booster::CheckAbort(const BatchCtrl& hc) {
  if (!hc.abort) return;
  throw std::runtime_error(StrFormat("Data-parallel training aborted on rank %d: another rank failed or did not arrive within %d s "
                                     "(listen_time_out of LGBM_NetworkInit); this booster cannot continue", net_->rank(), net_->timeout_s()));
}

void Booster::AbortPeers() {
  if (peers_.world == 0 || device_ < 0) return;
  cudaSetDevice(device_);
  k_peer_abort<<<1, 32, 0, copy_stream_>>>(peers_);
  cudaStreamSynchronize(copy_stream_);
}

// BoostFromScore of l1 / quantile / mape: PercentileFun / WeightedPercentileFun (upstream regression_objective.hpp) over this
// machine's labels, on the host (once per fit).
namespace {
template <typename T>
T HostPercentile(std::vector<T> data, double alpha, bool alpha_is_float) {
  const int cnt = static_cast<int>(data.size());
  if (cnt <= 1) return data[0];
  const double float_pos = alpha_is_float ? static_cast<double>((1.0f - static_cast<float>(alpha)) * cnt) : (1.0f - alpha) * cnt;
  const int pos = static_cast<int>(float_pos);
  if (pos < 1) return *std::max_element(data.begin(), data.end());
  if (pos >= cnt) return *std::min_element(data.begin(), data.end());
  const double bias = float_pos - pos;
  std::nth_element(data.begin(), data.begin() + (pos - 1), data.end(), std::greater<T>());      // [pos - 1]: the pos-th largest
  const T v1 = data[pos - 1];
  const T v2 = *std::max_element(data.begin() + pos, data.end());
  return static_cast<T>(v1 - (v1 - v2) * bias);
}
template <typename T1, typename T2>
T1 HostWeightedPercentile(const std::vector<T1>& data, const std::vector<T2>& weight, double alpha, bool alpha_is_float) {
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
}  // namespace

double Booster::PercentileInitScore() const {
  std::vector<float> lab(ds_->label.begin(), ds_->label.begin() + n_);
  const double a = obj_ == kQuantile ? cfg_.alpha : 0.5;
  const bool af = obj_ == kQuantile;
  if (obj_ == kMape) return static_cast<double>(HostWeightedPercentile<float, float>(lab, label_weight_, 0.5, false));
  if (has_weights_) return static_cast<double>(HostWeightedPercentile<float, float>(lab, std::vector<float>(ds_->weight.begin(), ds_->weight.begin() + n_), a, af));
  return static_cast<double>(HostPercentile<float>(lab, a, af));
}

// SerialTreeLearner::RenewTreeOutput for the finished tree (device: batch_tree.cuh "leaf re-fit"); with several machines the
// outputs of the machines that hold rows of a leaf are averaged (Network::GlobalSum of outputs and of the has-rows flags).
void Booster::RenewTreeOutput(HostTree* tree) {
  const int L = cfg_.num_leaves;
  const int nl = tree->num_leaves;
  if (nl <= 1) return;
  const bool weighted = has_weights_ || obj_ == kMape;
  const float* w = obj_ == kMape ? d_label_weight_.p : has_weights_ ? ds_->d_weight.p : nullptr;
  d_renew_off_.Reserve(L + 1);
  d_renew_out_.Reserve(2 * static_cast<size_t>(L));
  for (int k = 0; k < 2; ++k) { d_renew_res_[k].Reserve(n_); if (weighted) d_renew_w_[k].Reserve(n_); }
  k_renew_offsets<<<1, 32, 0, stream_>>>(d_nodes_, d_avail_.p, d_ctrl_, L, d_renew_off_.p);
  const dim3 cgrid(nl, std::max(1, std::min(64, static_cast<int>(n_ / (1024 * static_cast<int64_t>(nl)) + 1))));
  const int rf_pred_on = boost_ == kRF ? 1 : 0;
  const double rf_pred = boost_ == kRF ? rf_init_scores_[cls_] : 0.0;
  if (idx_es_ == 2)
    k_renew_collect<2><<<cgrid, 256, 0, stream_>>>(d_nodes_, d_avail_.p, d_ctrl_, d_renew_off_.p, d_idx_[0].p, d_idx_[1].p, ds_->d_label.p,
                                                   d_score_.p + cls_ * n_, w, d_renew_res_[0].p, d_renew_w_[0].p, rf_pred_on, rf_pred);
  else
    k_renew_collect<4><<<cgrid, 256, 0, stream_>>>(d_nodes_, d_avail_.p, d_ctrl_, d_renew_off_.p, d_idx_[0].p, d_idx_[1].p, ds_->d_label.p,
                                                   d_score_.p + cls_ * n_, w, d_renew_res_[0].p, d_renew_w_[0].p, rf_pred_on, rf_pred);
  size_t tmp = 0;
  const int n = static_cast<int>(n_);
  if (weighted)
    LGB_CUDA(cub::DeviceSegmentedRadixSort::SortPairs(nullptr, tmp, d_renew_res_[0].p, d_renew_res_[1].p, d_renew_w_[0].p, d_renew_w_[1].p, n, nl,
                                                      d_renew_off_.p, d_renew_off_.p + 1, 0, 64, stream_));
  else
    LGB_CUDA(cub::DeviceSegmentedRadixSort::SortKeys(nullptr, tmp, d_renew_res_[0].p, d_renew_res_[1].p, n, nl, d_renew_off_.p,
                                                     d_renew_off_.p + 1, 0, 64, stream_));
  d_cub_tmp_.Reserve(tmp);
  tmp = d_cub_tmp_.n;
  if (weighted)
    LGB_CUDA(cub::DeviceSegmentedRadixSort::SortPairs(d_cub_tmp_.p, tmp, d_renew_res_[0].p, d_renew_res_[1].p, d_renew_w_[0].p, d_renew_w_[1].p, n, nl,
                                                      d_renew_off_.p, d_renew_off_.p + 1, 0, 64, stream_));
  else
    LGB_CUDA(cub::DeviceSegmentedRadixSort::SortKeys(d_cub_tmp_.p, tmp, d_renew_res_[0].p, d_renew_res_[1].p, n, nl, d_renew_off_.p,
                                                     d_renew_off_.p + 1, 0, 64, stream_));
  k_renew_pick<<<CeilDiv(L, 64), 64, 0, stream_>>>(d_ctrl_, d_renew_off_.p, d_renew_res_[1].p, weighted ? d_renew_w_[1].p : nullptr,
                                                   obj_ == kQuantile ? cfg_.alpha : 0.5, obj_ == kQuantile ? 1 : 0, tree_dev_.leaf_value,
                                                   d_renew_out_.p, L);
  std::vector<double> out(2 * static_cast<size_t>(L));
  LGB_CUDA(cudaMemcpyAsync(out.data(), d_renew_out_.p, sizeof(double) * 2 * L, cudaMemcpyDeviceToHost, stream_));
  LGB_CUDA(cudaStreamSynchronize(stream_));
  LGB_CUDA(cudaGetLastError());
  launches_ += 4;
  if (net_->active()) {
    net_->HostAllReduceSum(out.data(), 2 * L);
    for (int l = 0; l < nl; ++l) out[l] = out[L + l] > 0.0 ? out[l] / out[L + l] : 0.0;
    LGB_CUDA(cudaMemcpyAsync(d_renew_out_.p, out.data(), sizeof(double) * nl, cudaMemcpyHostToDevice, stream_));
    k_set_leaf_values<<<CeilDiv(nl, 64), 64, 0, stream_>>>(tree_dev_.leaf_value, d_renew_out_.p, nl);
    LGB_CUDA(cudaStreamSynchronize(stream_));
  }
  for (int l = 0; l < nl; ++l) tree->leaf_value[l] = out[l];
}

// Copies the control block and the tree arrays back; false when the tree is not finished yet (more rounds are needed).
bool Booster::FetchTree(HostTree* t) {
  const int L = cfg_.num_leaves;
  // the copies wait for the tree (the event recorded behind its last round), not for the apply-tree pass enqueued after it: the
  // host decodes the tree and enqueues the next iteration while that pass runs
  BatchCtrl hc;
  LGB_CUDA(cudaStreamWaitEvent(copy_stream_, ev_rounds_done_, 0));
  LGB_CUDA(cudaMemcpyAsync(h_tree_blob_.data(), d_tree_blob_.p, tree_blob_bytes_, cudaMemcpyDeviceToHost, copy_stream_));
  LGB_CUDA(cudaMemcpyAsync(&hc, d_ctrl_, sizeof(BatchCtrl), cudaMemcpyDeviceToHost, copy_stream_));
  LGB_CUDA(cudaStreamSynchronize(copy_stream_));
  CheckAbort(hc);
  if (!hc.done) return false;
  last_rounds_ = hc.rounds;
  last_node_draws_ = hc.node_draws; last_searched_ = hc.last_searched;
  part_rows_total_ += hc.part_rows;
  rounds_total_ += hc.rounds; expansions_total_ += (hc.num_nodes - 1) / 2; splits_total_ += hc.num_leaves - 1;
  last_rows_visited_ += hc.hist_rows;
  hist_launches_ += hc.hist_launches;
  LGB_CUDA(cudaMemsetAsync(&d_ctrl_->hist_rows, 0, 3 * sizeof(long long), stream_));
  const int nl = hc.num_leaves;
  t->num_leaves = nl;
  t->shrinkage = 1.0;
  const unsigned char* b = h_tree_blob_.data();
  const double* dv = reinterpret_cast<const double*>(b);
  const size_t dbl = 4 * sizeof(double) * L, ints = 8 * sizeof(int) * L, floats = RoundUp<size_t>(sizeof(float) * L, 8);
  const int* ip = reinterpret_cast<const int*>(b + dbl);
  const float* fp = reinterpret_cast<const float*>(b + dbl + ints);
  const signed char* cp = reinterpret_cast<const signed char*>(b + dbl + ints + floats);
  const int ni = std::max(nl - 1, 0);
  t->leaf_value.assign(dv, dv + nl); t->leaf_weight.assign(dv + L, dv + L + nl);
  t->internal_value.assign(dv + 2 * L, dv + 2 * L + ni); t->internal_weight.assign(dv + 3 * L, dv + 3 * L + ni);
  t->left_child.assign(ip, ip + ni); t->right_child.assign(ip + L, ip + L + ni);
  t->split_feature_inner.assign(ip + 2 * L, ip + 2 * L + ni); t->threshold_in_bin.assign(ip + 3 * L, ip + 3 * L + ni);
  t->leaf_count.assign(ip + 4 * L, ip + 4 * L + nl); t->internal_count.assign(ip + 5 * L, ip + 5 * L + ni);
  t->leaf_depth.assign(ip + 7 * L, ip + 7 * L + nl);
  t->split_gain.assign(fp, fp + ni);
  t->decision_type.assign(cp, cp + ni);
  t->split_feature.resize(ni); t->threshold.resize(ni);
  t->num_cat = 0; t->cat_boundaries.assign(1, 0); t->cat_threshold.clear(); t->cat_bits_inner.clear();
  const unsigned int* cb = ds_->has_categorical ? reinterpret_cast<const unsigned int*>(b + tree_cat_off_) : nullptr;
  for (int i = 0; i < ni; ++i) {
    const int inner = t->split_feature_inner[i];
    const BinMapperH& m = ds_->mappers[ds_->used[inner]];
    t->split_feature[i] = ds_->used[inner];
    if (t->decision_type[i] & 1) {
      // Tree::SplitCategorical: the node's threshold is its index among the tree's categorical splits; the bins of the split become
      // a bitset over their category values (SerialTreeLearner::SplitInner: RealThreshold of every bin, Common::ConstructBitset)
      LGB_CHECK(cb != nullptr, "categorical node in a tree of a dataset without categorical features");
      std::vector<uint32_t> bits;
      for (int bin = 1; bin < m.num_bin; ++bin) {
        if (!((cb[i * 8 + (bin >> 5)] >> (bin & 31)) & 1u)) continue;
        const int cat = m.bin_2_cat[bin];
        if (static_cast<int>(bits.size()) < cat / 32 + 1) bits.resize(cat / 32 + 1, 0u);
        bits[cat / 32] |= 1u << (cat % 32);
      }
      t->threshold[i] = t->num_cat;
      t->threshold_in_bin[i] = t->num_cat;
      ++t->num_cat;
      t->cat_boundaries.push_back(t->cat_boundaries.back() + static_cast<int>(bits.size()));
      t->cat_threshold.insert(t->cat_threshold.end(), bits.begin(), bits.end());
    } else if (m.num_bin > kMaxBinsPerFeature) {
      // a wide feature's split arrives in column form (column of the feature, byte threshold): back to the logical bin
      const int c = inner - ds_->real_to_inner[ds_->used[inner]];
      t->threshold[i] = m.upper[c * kWideLocalBins + t->threshold_in_bin[i] - 1];
    } else {
      t->threshold[i] = m.upper[t->threshold_in_bin[i]];   // RealThreshold
    }
  }
  if (t->num_cat > 0) t->cat_bits_inner.assign(cb, cb + 8 * ni);
  return true;
}

// Adds the finished tree to the training scores.  For the row-wise objectives the same pass also produces the gradients of
// the next iteration (into the other gradient buffer) and their absolute maxima, which the next UpdateOneIter then picks up
// instead of running the gradient and max kernels; if no further iteration follows, they are simply never used.
void Booster::UpdateScores() {
  const int n = static_cast<int>(n_);
  const BinLayout& lay = ds_->layout;
  const int obj = (fuse_grad_ && obj_ == kBinary) ? 1 : (fuse_grad_ && obj_ == kL2) ? 2 : 0;
  const float* w = has_weights_ ? ds_->d_weight.p : nullptr;
  const int grid = std::min(CeilDiv(n, kApplyThreads), kNumSms * 8);
  float* g = d_grad_[grad_cur_ ^ 1].p;
  float* h = d_hess_[grad_cur_ ^ 1].p;
  if (obj != 0) LGB_CUDA(cudaMemsetAsync(&d_ctrl_->max_abs_g_bits, 0, 2 * sizeof(unsigned int), stream_));
#define APPLY_ARGS ds_->d_rm.p, static_cast<unsigned>(lay.rm_stride), d_score_.p + cls_ * n_, ds_->d_label.p, w, g, h, tree_dev_, ds_->d_fmeta.p, d_ctrl_, \
                   shrinkage_rate_, n, cfg_.sigmoid, label_weights_[0], label_weights_[1], mix_, 0
#define APPLY_C(OBJ, CAT)                                                                                            \
  do {                                                                                                               \
    if (lay.rm_stride == 32) k_bt_apply_tree<OBJ, 8, CAT><<<grid, kApplyThreads, 0, stream_>>>(APPLY_ARGS);          \
    else if (lay.rm_stride == 64) k_bt_apply_tree<OBJ, 16, CAT><<<grid, kApplyThreads, 0, stream_>>>(APPLY_ARGS);    \
    else k_bt_apply_tree<OBJ, 0, CAT><<<grid, kApplyThreads, 0, stream_>>>(APPLY_ARGS);                              \
  } while (0)
#define APPLY(OBJ) do { if (ds_->has_categorical) APPLY_C(OBJ, true); else APPLY_C(OBJ, false); } while (0)
  if (obj == 1) APPLY(1); else if (obj == 2) APPLY(2); else APPLY(0);
#undef APPLY_C
#undef APPLY
#undef APPLY_ARGS
  ApplyTreeToValidSets();
  k_bt_mark_applied<<<1, 1, 0, stream_>>>(d_ctrl_);
  LGB_CUDA(cudaGetLastError());
  grad_ready_ = obj != 0;
  launches_ += 2;
}

// Validation scores advance with the training scores: the same tree walk over the validation set's row records (scores only).
void Booster::ApplyTreeToValidSets() {
  for (ValidSet& v : valids_) {
    const Dataset* d = v.ds;
    const int n = static_cast<int>(d->num_data);
    if (n == 0) continue;
    const BinLayout& lay = d->layout;
    const int grid = std::min(CeilDiv(n, kApplyThreads), kNumSms * 8);
#define APPLY_V(WPR, CAT) k_bt_apply_tree<0, WPR, CAT><<<grid, kApplyThreads, 0, stream_>>>(d->d_rm.p, static_cast<unsigned>(lay.rm_stride), v.score.p + cls_ * d->num_data, \
        nullptr, nullptr, nullptr, nullptr, tree_dev_, ds_->d_fmeta.p, d_ctrl_, shrinkage_rate_, n, cfg_.sigmoid, 1.0, 1.0, mix_, 0)
    if (ds_->has_categorical) { if (lay.rm_stride == 32) APPLY_V(8, true); else if (lay.rm_stride == 64) APPLY_V(16, true); else APPLY_V(0, true); }
    else if (lay.rm_stride == 32) APPLY_V(8, false); else if (lay.rm_stride == 64) APPLY_V(16, false); else APPLY_V(0, false);
#undef APPLY_V
    ++launches_;
  }
}

void Booster::AddConstantToValidSets(double v) {
  for (ValidSet& s : valids_) {
    const int n = static_cast<int>(s.ds->num_data);
    if (n == 0) continue;
    k_add_constant<<<std::min(CeilDiv(n, 256), kNumSms * 8), 256, 0, stream_>>>(s.score.p + cls_ * s.ds->num_data, n, v);
    ++launches_;
  }
}

// ---------------------------------------------------------------------------------------------------- GBDT
bool Booster::UpdateOneIter() {
  LGB_CHECK(obj_ != kCustom, "objective=custom has no gradients of its own: call LGBM_BoosterUpdateOneIterCustom");
  return TrainOneIter(nullptr, nullptr);
}

// GBDT::TrainOneIter(gradients, hessians) with caller-supplied gradients (upstream src/boosting/gbdt.cpp): no automatic initial
// score, the arrays replace the objective's for this iteration, everything behind them is the same path.
bool Booster::UpdateOneIterCustom(const float* grad, const float* hess) {
  LGB_CHECK(grad != nullptr && hess != nullptr, "LGBM_BoosterUpdateOneIterCustom: grad / hess is null");
  return TrainOneIter(grad, hess);
}

bool Booster::TrainOneIter(const float* custom_grad, const float* custom_hess) {
  LGB_CHECK(!predict_only_, "this booster was loaded from a model string and can only predict");
  LGB_CUDA(cudaSetDevice(device_));
  const double kEps = 1e-15f;
  const bool custom = custom_grad != nullptr;
  LGB_CHECK(!(custom && K_ > 1 && obj_ != kCustom), "LGBM_BoosterUpdateOneIterCustom next to a multiclass objective is not supported (use objective=custom)");
  // ObtainAutomaticInitialScore (upstream src/boosting/gbdt.cpp): the objective's BoostFromScore over the labels of all machines,
  // then the mean over machines of what are identical values
  auto automatic_init_score = [&]() -> double {
    if (obj_ == kLambdarank || obj_ == kXendcg || obj_ == kCustom) return 0.0;
    double v;
    if (obj_ == kMulticlass) {
      v = std::log(std::max<double>(kEps, class_init_probs_[cls_]));
    } else if (obj_ == kMulticlassOva) {      // BinaryLogloss::BoostFromScore of the class's one-vs-rest problem
      double pavg = ova_suml_[cls_] / ova_sumw_;
      pavg = std::min(pavg, 1.0 - kEps);
      pavg = std::max<double>(pavg, kEps);
      v = std::log(pavg / (1.0f - pavg)) / cfg_.sigmoid;
    } else if (RenewsLeaves()) {
      // the (weighted) percentile of this machine's labels; ObtainAutomaticInitialScore then takes the mean over machines
      double acc = PercentileInitScore();
      net_->HostAllReduceSum(&acc, 1);
      v = acc / net_->world();
    } else {
      double sums[2];
      BoostFromScoreLocalSums(&sums[0], &sums[1]);
      net_->HostAllReduceSum(sums, 2);
      if (obj_ == kBinary || obj_ == kXentropy) {      // (cross_entropy: the mean label, no sigmoid parameter)
        double pavg = sums[0] / sums[1];
        pavg = std::min(pavg, 1.0 - kEps);
        pavg = std::max<double>(pavg, kEps);
        v = std::log(pavg / (1.0f - pavg)) / (obj_ == kBinary ? cfg_.sigmoid : 1.0);
      } else {
        v = sums[0] / sums[1];
        if (LogLink()) v = v > 0.0 ? std::log(v) : -std::numeric_limits<double>::infinity();      // Common::SafeLog of the mean label
      }
    }
    if (net_->active()) {
      double acc = 0.0;
      for (int s = 0; s < net_->world(); ++s) acc += v;
      v = acc / net_->world();
    }
    return v;
  };
  auto add_to_scores = [&](double v) {      // to the scores of class cls_
    k_add_constant<<<std::min(CeilDiv(static_cast<int>(n_), 256), kNumSms * 8), 256, 0, stream_>>>(d_score_.p + cls_ * n_, static_cast<int>(n_), v);
    AddConstantToValidSets(v);
    ++launches_;
  };
  const long long iter_total = iter_ + num_init_iteration_;
  if (boost_ == kRF) {
    LGB_CHECK(!custom, "LGBM_BoosterUpdateOneIterCustom with boosting=rf is not supported by this library");
    // RF::Boosting: the automatic initial scores are not added to the scores (every tree carries them); the gradients are taken
    // once, at those constant scores
    if (rf_init_scores_.empty()) {
      rf_init_scores_.assign(K_, 0.0);
      for (cls_ = 0; cls_ < K_; ++cls_) {
        if (!ds_->has_init_score && (cfg_.boost_from_average || F_ == 0)) {
          const double v = automatic_init_score();
          if (std::fabs(v) > kEps) rf_init_scores_[cls_] = v;
        }
      }
      cls_ = 0;
      DevBuf<double> tmp;
      tmp.Alloc(static_cast<size_t>(n_) * K_);
      for (int k = 0; k < K_; ++k)
        k_fill_double<<<std::min(CeilDiv(static_cast<int>(n_), 256), kNumSms * 8), 256, 0, stream_>>>(tmp.p + static_cast<size_t>(k) * n_, n_, rf_init_scores_[k]);
      ComputeGradients(tmp.p);
      LGB_CUDA(cudaStreamSynchronize(stream_));
    }
  }
  // GBDT::BoostFromAverage, once per class
  std::vector<double> init_scores(K_, 0.0);
  if (boost_ != kRF && !custom && models_.empty() && !ds_->has_init_score && (cfg_.boost_from_average || F_ == 0)) {
    for (cls_ = 0; cls_ < K_; ++cls_) {
      const double v = automatic_init_score();
      if (std::fabs(v) > kEps) {
        init_scores[cls_] = v;
        add_to_scores(v);
        LogInfo("Start training from score %lf", v);
      }
    }
    cls_ = 0;
  }
  if (boost_ == kDART) DroppingTrees();      // DART::GetTrainingScore: before the gradients are taken
  // Boosting(): the gradients of all classes from the scores before this iteration's trees
  const bool any_train = K_ > 1 ? std::any_of(class_need_train_.begin(), class_need_train_.end(), [](char c) { return c != 0; }) : need_train_;
  const bool fused = grad_ready_ && !custom;       // produced by the previous iteration's apply-tree pass
  if (any_train && F_ > 0) {
    if (fused) grad_cur_ ^= 1;
    else if (custom) {       // pageable sources are staged before the calls return
      LGB_CUDA(cudaMemcpyAsync(d_grad_[grad_cur_].p, custom_grad, sizeof(float) * n_ * K_, cudaMemcpyHostToDevice, stream_));
      LGB_CUDA(cudaMemcpyAsync(d_hess_[grad_cur_].p, custom_hess, sizeof(float) * n_ * K_, cudaMemcpyHostToDevice, stream_));
    } else if (boost_ != kRF) { ProfBegin("gradients"); ComputeGradients(); ProfEnd(); }
    grad_ready_ = false;
    Bagging();
  }
  bool should_continue = false;
  for (cls_ = 0; cls_ < K_; ++cls_) {
    HostTree tree;
    tree.num_leaves = 1;
    tree.leaf_value.assign(1, 0.0);
    const bool train_this = K_ > 1 ? class_need_train_[cls_] != 0 : need_train_;
    if (train_this && F_ > 0) {
      if (boost_ == kRF)      // the scores keep the running mean of the trees; the tree enters with the initial score as its bias
        mix_ = ScoreMix{static_cast<double>(iter_total), 1.0 / (iter_total + 1), std::fabs(rf_init_scores_[cls_]) > kEps ? rf_init_scores_[cls_] : 0.0, 0};
      ProfBegin("quantize"); PrepareGradients(fused && boost_ != kGOSS); ProfEnd();      // GOSS has rescaled gradients: new maxima
      TrainTree(&tree);
      if (tree.num_leaves <= 1) grad_ready_ = false;      // no tree, so the apply pass did nothing
    }
    if (boost_ == kRF) {
      if (tree.num_leaves > 1) {
        if (std::fabs(rf_init_scores_[cls_]) > kEps) tree.AddBias(rf_init_scores_[cls_]);
      } else if (static_cast<int>(models_.size()) < K_) {
        double output = rf_init_scores_[cls_];
        if (!train_this) output = automatic_init_score();      // objective->BoostFromScore(class)
        tree.AsConstantTree(output);
        const int grid = std::min(CeilDiv(static_cast<int>(n_), 256), kNumSms * 8);
        k_mix_constant<<<grid, 256, 0, stream_>>>(d_score_.p + cls_ * n_, n_, static_cast<double>(iter_total), output, 1.0 / (iter_total + 1));
        for (ValidSet& vs : valids_)
          k_mix_constant<<<grid, 256, 0, stream_>>>(vs.score.p + cls_ * vs.ds->num_data, vs.ds->num_data, static_cast<double>(iter_total), output, 1.0 / (iter_total + 1));
      }
      should_continue = true;      // a forest never stops early
    } else if (tree.num_leaves > 1) {
      should_continue = true;      // (the scores were advanced inside TrainTree, behind the tree's last round)
      tree.Shrinkage(shrinkage_rate_);
      if (std::fabs(init_scores[cls_]) > kEps) tree.AddBias(init_scores[cls_]);
    } else if (static_cast<int>(models_.size()) < K_) {
      // "only add default score one-time" (GBDT::TrainOneIter): with boost_from_average the initial score is already in the scores
      // (above) and only names the constant tree; without it — and with an objective and no init_score — it is obtained here
      if (!custom && obj_ != kCustom && !cfg_.boost_from_average && !ds_->has_init_score) {
        init_scores[cls_] = automatic_init_score();
        add_to_scores(init_scores[cls_]);
      }
      tree.AsConstantTree(init_scores[cls_]);
    }
    models_.push_back(std::move(tree));
  }
  cls_ = 0;
  // per-iteration histogram timing (only when enabled by the bench / tests)
  if (time_hist_ && hist_events_total_ > 0) {
    Sync();
    for (size_t i = 0; i < hist_events_total_; ++i) {
      float ms = 0;
      LGB_CUDA(cudaEventElapsedTime(&ms, hist_events_[i].first, hist_events_[i].second));
      last_hist_ms_ += ms;
    }
  }
  hist_events_total_ = 0;
  ++model_version_;
  if (!should_continue) {
    LogWarning("Stopped training because there are no more leaves that meet the split requirements");
    if (static_cast<int>(models_.size()) > K_) for (int k = 0; k < K_; ++k) models_.pop_back();
    ++model_version_;
    Sync();
    return true;
  }
  ++iter_;
  if (boost_ == kDART) {      // DART::TrainOneIter, after GBDT::TrainOneIter
    NormalizeDart();
    if (!cfg_.uniform_drop) { tree_weight_.push_back(shrinkage_rate_); sum_weight_ += shrinkage_rate_; }
  }
  return false;
}

ModelHeader Booster::MakeHeader() const {
  if (predict_only_) return loaded_header_;
  ModelHeader h;
  h.max_feature_idx = ds_->num_total_features - 1;
  if (obj_ == kBinary) h.objective = "binary sigmoid:" + FormatG(cfg_.sigmoid);
  else if (obj_ == kL2) h.objective = "regression";
  else if (obj_ == kMulticlass) { h.objective = "multiclass num_class:" + std::to_string(K_); h.num_class = K_; h.num_tree_per_iteration = K_; }
  else if (obj_ == kMulticlassOva) {
    h.objective = "multiclassova num_class:" + std::to_string(K_) + " sigmoid:" + FormatG(cfg_.sigmoid);
    h.num_class = K_; h.num_tree_per_iteration = K_;
  }
  else if (obj_ >= kHuber) h.objective = cfg_.objective;      // huber / fair / poisson / gamma / tweedie / regression_l1 / quantile / mape: the name alone
  else if (obj_ == kCustom) { h.objective = ""; h.num_class = K_; h.num_tree_per_iteration = K_; }      // no objective line without an objective function
  else h.objective = "lambdarank";
  h.average_output = boost_ == kRF;
  h.feature_names = ds_->feature_names;
  for (int j = 0; j < ds_->num_total_features; ++j) {
    const BinMapperH& m = ds_->mappers[j];
    if (m.is_trivial) h.feature_infos.push_back("none");
    else if (m.bin_type != 0) {      // BinMapper::bin_info_string, categorical: the categories of the bins joined by ':'
      std::string v;
      for (size_t b = 0; b < m.bin_2_cat.size(); ++b) { if (b) v.push_back(':'); v += std::to_string(m.bin_2_cat[b]); }
      h.feature_infos.push_back(v);
    } else h.feature_infos.push_back("[" + FormatG17(m.min_val) + ":" + FormatG17(m.max_val) + "]");
  }
  for (auto& kv : cfg_.raw) h.parameters.emplace_back(kv.first, kv.second);
  return h;
}

std::string Booster::SaveModelToString(int start_iteration, int num_iteration) const {
  return ModelToString(MakeHeader(), models_, start_iteration, num_iteration);
}

void Booster::CopyScores(double* out) const {
  LGB_CUDA(cudaSetDevice(device_));
  Sync();
  LGB_CUDA(cudaMemcpy(out, d_score_.p, sizeof(double) * n_ * K_, cudaMemcpyDeviceToHost));      // class-major [K][n]
}

int64_t Booster::NumValidRows(int i) const {
  LGB_CHECK(i >= 0 && i < static_cast<int>(valids_.size()), "validation data %d does not exist", i + 1);
  return valids_[i].ds->num_data;
}

int64_t Booster::CopyValidScores(int i, double* out) const {
  const int64_t n = NumValidRows(i);
  LGB_CUDA(cudaSetDevice(device_));
  Sync();
  LGB_CUDA(cudaMemcpy(out, valids_[i].score.p, sizeof(double) * n * K_, cudaMemcpyDeviceToHost));
  return n;
}

void Booster::CopyGradients(float* g, float* h) const {
  LGB_CUDA(cudaSetDevice(device_));
  Sync();
  LGB_CUDA(cudaMemcpy(g, d_grad_[grad_cur_].p, sizeof(float) * n_ * K_, cudaMemcpyDeviceToHost));
  LGB_CUDA(cudaMemcpy(h, d_hess_[grad_cur_].p, sizeof(float) * n_ * K_, cudaMemcpyDeviceToHost));
}

void Booster::CopyQuantGradients(int8_t* g, int8_t* h, double* scales) const {
  LGB_CHECK(lp_.quantized, "quantized gradients exist only with use_quantized_grad=true");
  LGB_CUDA(cudaSetDevice(device_));
  Sync();
  std::vector<uint16_t> tmp(n_);
  LGB_CUDA(cudaMemcpy(tmp.data(), d_gh16_.p, sizeof(uint16_t) * n_, cudaMemcpyDeviceToHost));
  for (int64_t i = 0; i < n_; ++i) { g[i] = static_cast<int8_t>(tmp[i] >> 8); h[i] = static_cast<int8_t>(tmp[i] & 0xff); }
  BatchCtrl hc;
  LGB_CUDA(cudaMemcpy(&hc, d_ctrl_, sizeof(BatchCtrl), cudaMemcpyDeviceToHost));
  scales[0] = hc.grad_scale; scales[1] = hc.hess_scale;
}

// ---------------------------------------------------------------------------------------------------- scoring
Booster::Booster(const std::string& model_str) : ds_(nullptr) {
  predict_only_ = true;
  ParseModelString(model_str, &loaded_header_, &models_);
  const std::string& o = loaded_header_.objective;
  if (o.rfind("binary", 0) == 0) {
    obj_ = kBinary;
    const size_t p = o.find("sigmoid:");
    if (p != std::string::npos) cfg_.sigmoid = strtod(o.c_str() + p + 8, nullptr);
  } else if (o.rfind("lambdarank", 0) == 0 || o.rfind("rank_xendcg", 0) == 0) {
    obj_ = kLambdarank;      // (scoring only: both ranking objectives leave the raw score as the prediction)
  } else if (o.rfind("multiclassova", 0) == 0) {
    obj_ = kMulticlassOva;
    K_ = loaded_header_.num_tree_per_iteration;
    LGB_CHECK(K_ >= 2 && K_ == loaded_header_.num_class, "Model string: multiclassova needs num_tree_per_iteration == num_class >= 2");
    const size_t p = o.find("sigmoid:");
    if (p != std::string::npos) cfg_.sigmoid = strtod(o.c_str() + p + 8, nullptr);
  } else if (o.rfind("multiclass", 0) == 0) {
    obj_ = kMulticlass;
    K_ = loaded_header_.num_tree_per_iteration;
    LGB_CHECK(K_ >= 2 && K_ == loaded_header_.num_class, "Model string: multiclass needs num_tree_per_iteration == num_class >= 2");
  } else if (o == "cross_entropy") {
    obj_ = kXentropy;
    cfg_.objective = o;
  } else if (o == "huber" || o == "fair" || o == "poisson" || o == "gamma" || o == "tweedie") {
    obj_ = o == "huber" ? kHuber : o == "fair" ? kFair : o == "poisson" ? kPoisson : o == "gamma" ? kGamma : kTweedie;
    cfg_.objective = o;
  } else if (o == "regression_l1" || o == "quantile" || o == "mape") {
    obj_ = o == "regression_l1" ? kL1 : o == "quantile" ? kQuantile : kMape;      // identity output, as plain regression
    cfg_.objective = o;
  } else if (o.empty() || o.rfind("custom", 0) == 0) {
    obj_ = kCustom;      // trained with caller-supplied gradients: predictions are raw scores
    K_ = std::max(1, loaded_header_.num_tree_per_iteration);
  } else {
    LGB_CHECK(o.rfind("regression", 0) == 0, "Model string: objective '%s' is not supported by this library", o.c_str());
    obj_ = kL2;
  }
  for (auto& kv : loaded_header_.parameters) cfg_.raw[kv.first] = kv.second;
  LGB_CHECK(loaded_header_.num_tree_per_iteration == K_ && loaded_header_.num_class == K_,
            "Model string: num_class=%d / num_tree_per_iteration=%d do not fit objective '%s'", loaded_header_.num_class,
            loaded_header_.num_tree_per_iteration, o.c_str());
  LGB_CHECK(models_.size() % K_ == 0, "Model string: %d trees are not a multiple of num_tree_per_iteration", static_cast<int>(models_.size()));
  if (loaded_header_.average_output) boost_ = kRF;
  device_ = -1;      // resolved at the first prediction: loading and re-saving a model needs no GPU
  iter_ = static_cast<int>(models_.size()) / K_;
  model_version_ = 1;
}

void Booster::MergeFrom(const Booster& other) {
  LGB_CHECK(!predict_only_, "LGBM_BoosterMerge: the receiving booster must be a training booster");
  LGB_CHECK(other.NumTotalFeatures() == NumTotalFeatures(), "LGBM_BoosterMerge: the merged model has %d features, the training data %d",
            other.NumTotalFeatures(), NumTotalFeatures());
  LGB_CHECK(other.K_ == K_, "LGBM_BoosterMerge: the merged model has %d trees per iteration, this booster %d", other.K_, K_);
  std::vector<HostTree> merged(other.models_);
  num_init_iteration_ = static_cast<int>(merged.size()) / K_;
  for (HostTree& t : models_) merged.push_back(std::move(t));
  models_ = std::move(merged);
  ++model_version_;
}

int Booster::NumTotalFeatures() const { return predict_only_ ? loaded_header_.max_feature_idx + 1 : ds_->num_total_features; }

void Booster::FeatureImportance(int num_iteration, int importance_type, double* out) const {
  LGB_CHECK(importance_type == 0 || importance_type == 1, "FeatureImportance: importance_type must be 0 (split) or 1 (gain)");
  const int nf = NumTotalFeatures();
  for (int j = 0; j < nf; ++j) out[j] = 0.0;
  const int n = num_iteration > 0 ? std::min<int>(num_iteration * K_, static_cast<int>(models_.size())) : static_cast<int>(models_.size());
  for (int t = 0; t < n; ++t) {
    const HostTree& tr = models_[t];
    for (int k = 0; k < tr.num_leaves - 1; ++k) {
      if (tr.split_gain[k] <= 0.0f) continue;      // upstream counts only splits with positive gain
      out[tr.split_feature[k]] += importance_type == 0 ? 1.0 : static_cast<double>(tr.split_gain[k]);
    }
  }
}

std::string Booster::DumpModel(int start_iteration, int num_iteration, int importance_type) const {
  std::vector<double> imp(std::max(NumTotalFeatures(), 1), 0.0);
  FeatureImportance(num_iteration, importance_type, imp.data());
  return ModelToJson(MakeHeader(), models_, start_iteration, num_iteration, imp);
}

void Booster::TreeRange(int start_iteration, int num_iteration, int* t0, int* t1) const {
  const int n = static_cast<int>(models_.size());      // iterations -> trees (K_ per iteration)
  *t0 = std::min(std::max(0, start_iteration) * K_, n);
  *t1 = num_iteration > 0 ? std::min(n, *t0 + num_iteration * K_) : n;
}

int64_t Booster::CalcNumPredict(int64_t nrow, int predict_type, int start_iteration, int num_iteration) const {
  int t0, t1;
  TreeRange(start_iteration, num_iteration, &t0, &t1);
  if (predict_type == kPredictContrib) return nrow * K_ * (NumTotalFeatures() + 1);
  return predict_type == kPredictLeaf ? nrow * (t1 - t0) : nrow * K_;
}

void Booster::UploadModel() {
  if (uploaded_version_ == model_version_) return;
  std::vector<int> node_off(1, 0), leaf_off(1, 0), feat, left, right, icount, lcount, depth;
  std::vector<double> thr, leaf, expected;
  std::vector<signed char> dt;
  std::vector<int> cat_begin, cat_words;      // per node: its category set in cat_thr (categorical nodes), else 0 / 0
  std::vector<unsigned int> cat_thr;
  model_max_depth_ = 0;
  for (const HostTree& t : models_) {
    const int ni = t.num_leaves - 1;
    const bool has_counts = static_cast<int>(t.internal_count.size()) == ni && static_cast<int>(t.leaf_count.size()) == t.num_leaves;
    const int cat_base = static_cast<int>(cat_thr.size());
    if (ni > 0) cat_thr.insert(cat_thr.end(), t.cat_threshold.begin(), t.cat_threshold.end());
    for (int i = 0; i < ni; ++i) {
      if (t.decision_type[i] & 1) {
        const int k = static_cast<int>(t.threshold[i]);
        cat_begin.push_back(cat_base + t.cat_boundaries[k]); cat_words.push_back(t.cat_boundaries[k + 1] - t.cat_boundaries[k]);
      } else { cat_begin.push_back(0); cat_words.push_back(0); }
      feat.push_back(t.split_feature[i]); thr.push_back(t.threshold[i]); dt.push_back(t.decision_type[i]);
      left.push_back(t.left_child[i]); right.push_back(t.right_child[i]);
      icount.push_back(has_counts ? t.internal_count[i] : 0);
    }
    for (int i = 0; i < t.num_leaves; ++i) { leaf.push_back(t.leaf_value[i]); lcount.push_back(has_counts ? t.leaf_count[i] : 0); }
    // Tree::ExpectedValue and the depth TreeSHAP's path storage is sized for
    double ev = t.leaf_value[0];
    int md = 0;
    if (ni > 0) {
      ev = 0.0;
      const double total = has_counts ? static_cast<double>(t.internal_count[0]) : 1.0;
      for (int i = 0; i < t.num_leaves; ++i) ev += ((has_counts ? t.leaf_count[i] : 0) / total) * t.leaf_value[i];
      std::vector<std::pair<int, int>> stack(1, {0, 0});
      while (!stack.empty()) {
        auto [nd, d] = stack.back();
        stack.pop_back();
        if (nd < 0) { md = std::max(md, d); continue; }
        stack.push_back({t.left_child[nd], d + 1});
        stack.push_back({t.right_child[nd], d + 1});
      }
    }
    expected.push_back(ev); depth.push_back(md);
    model_max_depth_ = std::max(model_max_depth_, md);
    node_off.push_back(static_cast<int>(feat.size()));
    leaf_off.push_back(static_cast<int>(leaf.size()));
  }
  auto up = [&](auto& buf, const auto& v) {
    buf.Alloc(std::max<size_t>(v.size(), 1));
    if (!v.empty()) LGB_CUDA(cudaMemcpyAsync(buf.p, v.data(), sizeof(v[0]) * v.size(), cudaMemcpyHostToDevice, stream_));
  };
  up(dp_node_off_, node_off); up(dp_leaf_off_, leaf_off); up(dp_feat_, feat); up(dp_left_, left); up(dp_right_, right);
  up(dp_cat_begin_, cat_begin); up(dp_cat_words_, cat_words); up(dp_cat_thr_, cat_thr);
  up(dp_thr_, thr); up(dp_leaf_, leaf); up(dp_dt_, dt); up(dp_icount_, icount); up(dp_lcount_, lcount); up(dp_expected_, expected); up(dp_depth_, depth);
  LGB_CUDA(cudaStreamSynchronize(stream_));      // the host vectors go out of scope
  uploaded_version_ = model_version_;
}

void Booster::PredictForMat(const void* data, int data_type, int64_t nrow, int ncol, bool row_major, int predict_type,
                            int start_iteration, int num_iteration, double* out, int64_t* out_len) {
  LGB_CHECK(data_type == 0 || data_type == 1, "Predict: data_type must be C_API_DTYPE_FLOAT32 or FLOAT64");
  LGB_CHECK(predict_type == kPredictNormal || predict_type == kPredictRaw || predict_type == kPredictLeaf || predict_type == kPredictContrib,
            "Predict: unknown predict_type %d", predict_type);
  LGB_CHECK(ncol == NumTotalFeatures(), "Predict: expected %d columns, got %d", NumTotalFeatures(), ncol);
  LGB_CHECK(nrow >= 0, "Predict: negative row count");
  std::lock_guard<std::mutex> lock(predict_mu_);
  if (device_ < 0) device_ = ResolveDevice(cfg_);
  LGB_CUDA(cudaSetDevice(device_));
  UploadModel();
  int t0, t1;
  TreeRange(start_iteration, num_iteration, &t0, &t1);
  const int per_row = predict_type == kPredictLeaf ? (t1 - t0) : predict_type == kPredictContrib ? K_ * (ncol + 1) : K_;
  if (out_len) *out_len = nrow * per_row;
  if (nrow == 0 || per_row == 0) return;
  PredModelDev m;
  m.node_off = dp_node_off_.p; m.leaf_off = dp_leaf_off_.p; m.split_feature = dp_feat_.p; m.threshold = dp_thr_.p;
  m.decision_type = dp_dt_.p; m.left = dp_left_.p; m.right = dp_right_.p; m.leaf_value = dp_leaf_.p;
  m.internal_count = dp_icount_.p; m.leaf_count = dp_lcount_.p; m.expected = dp_expected_.p; m.max_depth = dp_depth_.p;
  m.cat_begin = dp_cat_begin_.p; m.cat_words = dp_cat_words_.p; m.cat_threshold = dp_cat_thr_.p;
  // feature contributions: TreeSHAP recurses to the depth of the deepest tree and keeps (D + 1)(D + 2) / 2 path elements per thread
  const int shap_threads = kNumSms * 2 * 128;
  const int path_elems = (model_max_depth_ + 1) * (model_max_depth_ + 2) / 2;
  if (predict_type == kPredictContrib) {
    for (int t = t0; t < t1; ++t)
      LGB_CHECK(models_[t].num_leaves <= 1 || (static_cast<int>(models_[t].internal_count.size()) == models_[t].num_leaves - 1 && models_[t].internal_count[0] > 0),
                "feature contributions need the trees' data counts (internal_count / leaf_count), which tree %d lacks", t);
    dp_paths_.Reserve(static_cast<size_t>(shap_threads) * path_elems * sizeof(ShapPathElement));
    static bool stack_set = false;
    size_t need = 2048 + static_cast<size_t>(model_max_depth_ + 2) * 256;
    size_t cur = 0;
    LGB_CUDA(cudaDeviceGetLimit(&cur, cudaLimitStackSize));
    if (!stack_set || cur < need) { LGB_CUDA(cudaDeviceSetLimit(cudaLimitStackSize, std::max(cur, need))); stack_set = true; }
  }
  const size_t esz = data_type == 1 ? sizeof(double) : sizeof(float);
  // row-major input is processed in chunks of rows; column-major input in one piece
  const int64_t chunk_rows = row_major ? std::max<int64_t>(1, std::min<int64_t>(nrow, (int64_t(256) << 20) / (esz * ncol))) : nrow;
  dp_in_.Reserve(static_cast<size_t>(chunk_rows) * ncol * esz);
  dp_out_.Reserve(static_cast<size_t>(chunk_rows) * per_row);
  const int sig_on = (obj_ == kBinary || obj_ == kMulticlassOva || obj_ == kXentropy) ? 1 : LogLink() ? 2 : obj_ == kMulticlass ? 3 : 0;      // ConvertOutput: sigmoid, exp, softmax or none
  for (int64_t r0 = 0; r0 < nrow; r0 += chunk_rows) {
    const int64_t nr = std::min(chunk_rows, nrow - r0);
    const unsigned char* src = static_cast<const unsigned char*>(data) + (row_major ? static_cast<size_t>(r0) * ncol * esz : 0);
    LGB_CUDA(cudaMemcpyAsync(dp_in_.p, src, static_cast<size_t>(nr) * ncol * esz, cudaMemcpyHostToDevice, stream_));
    const long long rs = row_major ? ncol : 1, cs = row_major ? 1 : nr;
    const int grid = static_cast<int>(std::min<int64_t>(CeilDiv(static_cast<int>(std::min<int64_t>(nr, 1 << 30)), 128), kNumSms * 16));
    if (predict_type == kPredictContrib) {
      LGB_CUDA(cudaMemsetAsync(dp_out_.p, 0, sizeof(double) * nr * per_row, stream_));
      const int sgrid = std::min(grid, shap_threads / 128);
      const int avg = boost_ == kRF ? (t1 - t0) / K_ : 0;
      if (data_type == 1)
        k_predict_contrib<double><<<sgrid, 128, 0, stream_>>>(reinterpret_cast<const double*>(dp_in_.p), nr, rs, cs, m, t0, t1, K_, ncol, avg,
                                                              reinterpret_cast<ShapPathElement*>(dp_paths_.p), path_elems, dp_out_.p);
      else
        k_predict_contrib<float><<<sgrid, 128, 0, stream_>>>(reinterpret_cast<const float*>(dp_in_.p), nr, rs, cs, m, t0, t1, K_, ncol, avg,
                                                             reinterpret_cast<ShapPathElement*>(dp_paths_.p), path_elems, dp_out_.p);
    } else if (data_type == 1)
      k_predict<double><<<grid, 128, 0, stream_>>>(reinterpret_cast<const double*>(dp_in_.p), nr, rs, cs, m, t0, t1, K_, boost_ == kRF ? (t1 - t0) / K_ : 0, predict_type, sig_on,
                                                   obj_ == kXentropy ? 1.0 : cfg_.sigmoid, dp_out_.p);
    else
      k_predict<float><<<grid, 128, 0, stream_>>>(reinterpret_cast<const float*>(dp_in_.p), nr, rs, cs, m, t0, t1, K_, boost_ == kRF ? (t1 - t0) / K_ : 0, predict_type, sig_on,
                                                  obj_ == kXentropy ? 1.0 : cfg_.sigmoid, dp_out_.p);
    LGB_CUDA(cudaGetLastError());
    LGB_CUDA(cudaMemcpyAsync(out + static_cast<size_t>(r0) * per_row, dp_out_.p, sizeof(double) * nr * per_row, cudaMemcpyDeviceToHost, stream_));
    LGB_CUDA(cudaStreamSynchronize(stream_));
  }
}

void Booster::DebugRootHistogram(long long* out) {
  LGB_CUDA(cudaSetDevice(device_));
  k_bt_init<<<CeilDiv(cfg_.num_leaves, 256), 256, 0, stream_>>>(d_nodes_, d_avail_.p, d_ctrl_, tree_dev_, cfg_.num_leaves,
                                                                  static_cast<int>(n_), static_cast<int>(n_global_));
  LaunchHist(true);
  Sync();
  // out: [F][256][2] (sum g, sum h); the quantized work area is packed, one int64 per bin
  if (lp_.quantized) {
    std::vector<long long> pk(work_elems_);
    LGB_CUDA(cudaMemcpy(pk.data(), d_hist_work_.p + (exchange_seq_ & 1u) * work_stride_, sizeof(long long) * work_elems_, cudaMemcpyDeviceToHost));
    for (int i = 0; i < work_elems_; ++i) {
      const long long h = pk[i] & 0xffffffffLL;
      out[2 * i] = (pk[i] - h) >> 32;
      out[2 * i + 1] = h;
    }
  } else {
    LGB_CUDA(cudaMemcpy(out, d_hist_work_.p + (exchange_seq_ & 1u) * work_stride_, sizeof(long long) * work_elems_, cudaMemcpyDeviceToHost));
  }
}

void Booster::TimerStart() {
  LGB_CUDA(cudaSetDevice(device_));
  if (!timer_ev_[0]) { LGB_CUDA(cudaEventCreate(&timer_ev_[0])); LGB_CUDA(cudaEventCreate(&timer_ev_[1])); }
  LGB_CUDA(cudaEventRecord(timer_ev_[0], stream_));
}

float Booster::TimerStop() {
  LGB_CHECK(timer_ev_[0] != nullptr, "TimerStop without TimerStart");
  LGB_CUDA(cudaEventRecord(timer_ev_[1], stream_));
  LGB_CUDA(cudaEventSynchronize(timer_ev_[1]));
  float ms = 0;
  LGB_CUDA(cudaEventElapsedTime(&ms, timer_ev_[0], timer_ev_[1]));
  return ms;
}

float Booster::TimeRootHistogram(int repeats) {
  LGB_CUDA(cudaSetDevice(device_));
  k_bt_init<<<CeilDiv(cfg_.num_leaves, 256), 256, 0, stream_>>>(d_nodes_, d_avail_.p, d_ctrl_, tree_dev_, cfg_.num_leaves,
                                                                  static_cast<int>(n_), static_cast<int>(n_global_));
  const bool was = time_hist_;
  time_hist_ = true;
  float total = 0;
  for (int r = 0; r < repeats; ++r) {
    hist_events_total_ = 0;
    LaunchHist(true);
    Sync();
    float ms = 0;
    LGB_CUDA(cudaEventElapsedTime(&ms, hist_events_[0].first, hist_events_[0].second));
    total += ms;
  }
  time_hist_ = was;
  hist_events_total_ = 0;
  return total / repeats;
}

}  // namespace lgb
