// Batched best-first tree growth.
//
// Leaf-wise growth picks, one at a time, the leaf with the largest split gain — a chain of (num_leaves - 1) dependent steps,
// each made of several latency-bound launches (partition, histogram, split search; profiles/r01_*: ~75 us per step of which
// only a fraction is work).  The result of EXPANDING a leaf (partitioning its rows, building its children's histograms and
// finding their best splits) does not depend on when the leaf is split, so several leaves are expanded per round and the
// sequential order is recovered afterwards:
//   * select  — one CTA REPLAYS the reference order: while the best available leaf (gain desc, feature asc, leaf id asc) has
//               already been expanded, its split is COMMITTED (Tree::Split bookkeeping, leaf ids handed out in commit order,
//               exactly as the sequential learner does); replay stops at the first leaf that still has to be expanded, or
//               when the leaf budget / positive gains run out.  Then the next batch is chosen: the best not-yet-expanded
//               leaves whose rank among the available leaves is below the remaining budget (a leaf of rank r can only be
//               split after r - 1 others, so deeper ranks can never be committed).
//   * expand  — for every leaf of the batch at once: stable partition (K7), histogram of the smaller child (K4), sibling by
//               subtraction + split search for both children (K5/K6).
// The first leaf of every batch is always committed by the next replay, so a tree needs at most as many rounds as the
// sequential learner and typically 3-4x fewer; expansions that are never committed are simply dropped (their parent keeps
// its contiguous row range).  The committed tree is identical to the sequential learner's.
#pragma once
#include "kernels.cuh"

namespace lgb {

// Data-parallel learner, histogram exchange over peer memory (NVLink / NVSwitch) instead of an NCCL launch per round: every
// rank's work areas and mailbox (struct PeerMail / PeerTable: booster.h) are mapped into every other rank's address space (CUDA
// IPC).  The histogram kernel's last CTA stores "exchange #seq is complete" into every peer's mailbox; the split-search kernel
// polls its own (local) mailbox until every rank's word has reached seq and then sums the ranks' bins while it loads them —
// integer sums, so every rank obtains the same totals whatever the order.  Two forms:
//   * all-to-all (narrow data, a round is latency-bound): every rank sums every feature of every rank and searches all
//     features redundantly, so all ranks reach the same decision with ONE exchange per round;
//   * feature-sliced (wide data; the reduce-scatter by feature + all-gather of best splits of upstream's data-parallel
//     learner): rank g sums and searches only features [f_lo, f_hi) — 1/G of the remote reads and of the search — and stores
//     its best split per child into every peer's mailbox; the replay kernel takes the best of the G records.
// The work areas and split records are double-buffered by the parity of seq: a rank can be at most one exchange ahead of its
// slowest peer (its next search waits for that peer's next histogram), so the buffer it clears for exchange seq + 2 is no
// longer being read.  A wait is bounded: after PeerTable::timeout_ns, or when a peer has raised its abort word, the kernel
// sets BatchCtrl::abort and every later kernel of the fit returns at once; the host then fails the call.
__device__ __forceinline__ unsigned int LoadAcquireSys(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void StoreReleaseSys(unsigned int* p, unsigned int v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long GlobalTimerNs() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

struct NodeDev {
  int begin, count;        // local row range (positions in index buffer `buf`; buf < 0: identity, the root)
  int buf;
  int global_count;        // count used by the split finder and stored in the tree: true (serial) or estimated (data-parallel)
  long long int_g, int_h;
  double sum_g, sum_h;
  int depth, hist_slot;
  int state;               // 0: leaf whose children are not known; 1: expanded (children known, split not committed); 2: committed
  int left, right;         // child nodes once expanded
  int pad_;
  double cmin, cmax;       // monotone constraints: the output range of this leaf (BasicLeafConstraints entry)
  // what the replay needs to know about the children of an expanded node, written by the children's split search
  // ([0] = left child, [1] = right child), so that it never has to chase the child nodes
  int child_feat[2], child_count[2];
  double child_gain[2];
  SplitDev best;
};

// Everything the kernels of a round need to know about one expansion.  Written by the select step; the outcome of the
// partition (how many rows went left) is read from the item's cursor, so no kernel has to wait for a "publish" step:
// ChildGeom() derives the children's row ranges, which child is the smaller one and whether they are searched at all.
struct BatchItem {
  int node, left, right;                        // parent node and the node ids reserved for its children
  int feature, threshold, default_left, nan_bin;
  int src_begin, src_count, src_buf;            // the parent's rows
  int parent_slot, fresh_slot;                  // pool slots: the larger child takes over the parent's, the smaller gets the fresh one
  int child_depth;
  int search;                                   // 0: max_depth reached, the children are never searched
  int est_left, est_right;                      // the split's count estimates (the data-parallel learner's global counts)
  long long left_int_g, left_int_h, right_int_g, right_int_h;
  double left_sum_g, left_sum_h, right_sum_g, right_sum_h;
  int is_cat, pad_;                             // categorical split: rows whose bin has its bit set in cat_bits go left
  unsigned int cat_bits[8];
  // monotone constraints: the parent's output range, the split feature's direction and the two outputs (the children's ranges
  // follow from them: BasicLeafConstraints::Update)
  int mono, pad2_;
  double cmin, cmax, left_output, right_output;
};

struct BatchCtrl {
  int num_leaves;          // leaves of the committed tree
  int done;
  int num_nodes;           // nodes handed out so far
  int num_slots;           // histogram pool slots handed out so far
  int batch_n;
  int spec_used;           // expansions beyond the first of each batch so far (statistics)
  int rounds;
  int applied;             // the finished tree has been added to the scores
  int abort;               // a wait for a peer rank failed: every kernel returns at once, the host raises the error
  // feature_fraction_bynode: how many by-node column samples the searches of this tree have used so far (the next searched
  // expansion's children take samples node_draws and node_draws + 1), and whether the last expansion's children were searched
  int node_draws, last_searched, pad0_;
  unsigned long long part_cursor[kMaxBatch];   // per batch item: (lefts placed << 32) | rights placed
  BatchItem batch[kMaxBatch];
  int find_count[kMaxBatch][2];
  unsigned int max_abs_g_bits, max_abs_h_bits;
  double grad_scale, hess_scale, inv_grad_scale, inv_hess_scale;
  long long hist_rows, hist_launches, part_rows;      // statistics, cleared by the host after every tree
  // optional timeline for experiments (LGB_B200_TRACE=1): (event id, %globaltimer ns) pairs written by CTA 0 of every kernel
  long long* trace;
  int trace_n, trace_cap;
  // bagging: bit r = row r is in this iteration's bag (nullptr: every row is).  Rows outside the bag carry zero gradients into
  // the root histogram and are dropped by the root's partition, so below the root only rows of the bag exist.
  const unsigned int* bag_bits;
};

__constant__ int c_trace_on;      // set by the host while LGB_B200_TRACE records a tree
__device__ __forceinline__ void Trace(BatchCtrl* ctrl, int id) {
  if (!c_trace_on) return;      // a constant-bank read: the disabled case must not put a global load on thread 0's critical path
  if (ctrl->trace == nullptr) return;
  const unsigned long long t = GlobalTimerNs();
  const int i = atomicAdd(&ctrl->trace_n, 1);
  if (i < ctrl->trace_cap) { ctrl->trace[2 * i] = id; ctrl->trace[2 * i + 1] = static_cast<long long>(t); }
}
#define LGB_TRACE(ctrl, id) do { if (blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && threadIdx.x == 0) Trace(ctrl, id); } while (0)

__device__ __forceinline__ bool Finished(const BatchCtrl* ctrl) { return (ctrl->done | ctrl->abort) != 0; }

// Waits until *flag (a word of this rank's own mailbox, stored by a peer) has reached seq.  False: gave up (ctrl->abort is set).
__device__ __forceinline__ bool PeerWait(const unsigned int* flag, unsigned int seq, const PeerTable& pt, BatchCtrl* ctrl) {
  unsigned long long t0 = 0;
  for (unsigned int n = 1;; ++n) {
    if (static_cast<int>(LoadAcquireSys(flag) - seq) >= 0) return true;
    if ((n & 255u) == 0) {
      const PeerMail* me = pt.mail[pt.rank];
      bool dead = *reinterpret_cast<volatile int*>(&ctrl->abort) != 0;
      for (int r = 0; r < pt.world; ++r) dead |= LoadAcquireSys(&me->abort[r]) != 0;
      const unsigned long long t = GlobalTimerNs();
      if (t0 == 0) t0 = t;
      if (dead || t - t0 > pt.timeout_ns) { ctrl->abort = 1; __threadfence(); return false; }
    }
  }
}

struct ChildGeom {
  int nl, nr;              // rows that went left / right on this GPU
  int rbegin;              // first position of the right child's rows (the rights are placed from the parent's back)
  int lc, rc;              // counts the split search uses: nl / nr (serial learner) or the estimates (data-parallel)
  int buf;                 // index buffer the children live in
  bool left_smaller, find;
};
__device__ __forceinline__ ChildGeom MakeGeom(const BatchItem& it, unsigned long long cursor, const LearnerParams& P) {
  ChildGeom g;
  g.nl = static_cast<int>(cursor >> 32);
  g.nr = static_cast<int>(cursor & 0xffffffffu);      // = src_count - nl except below a bagged root, which drops rows
  g.rbegin = it.src_begin + it.src_count - g.nr;
  g.lc = P.dp ? it.est_left : g.nl;
  g.rc = P.dp ? it.est_right : g.nr;
  g.buf = it.src_buf < 0 ? 0 : 1 - it.src_buf;
  g.left_smaller = g.lc < g.rc;
  g.find = it.search != 0 && !(g.rc < P.min_data_in_leaf * 2 && g.lc < P.min_data_in_leaf * 2);      // BeforeFindBestSplit
  return g;
}

// ------------------------------------------------------------------------------------------------ max |g|, max |h| over all ranks
// Quantisation scales need the maxima over ALL rows: lane r stores this rank's pair into rank r's mailbox and stamps it with the
// number of the gradient pass, then waits for rank r's pair in its own mailbox; the warp's maximum replaces the local one.
// (One warp, ~one NVLink store latency; replaces an NCCL all-reduce launch per iteration.)
__global__ void k_peer_max(BatchCtrl* ctrl, PeerTable pt, unsigned int pass) {
  if (ctrl->abort) return;
  const int lane = threadIdx.x;
  const unsigned int g = ctrl->max_abs_g_bits, h = ctrl->max_abs_h_bits;
  unsigned int mg = 0u, mh = 0u;
  if (lane < pt.world) {
    PeerMail* dst = pt.mail[lane];
    dst->max_vals[pass & 1u][pt.rank][0] = g;
    dst->max_vals[pass & 1u][pt.rank][1] = h;
    __threadfence_system();
    StoreReleaseSys(&dst->max_seq[pt.rank], pass);
    const PeerMail* me = pt.mail[pt.rank];
    if (PeerWait(&me->max_seq[lane], pass, pt, ctrl)) {
      mg = *reinterpret_cast<const volatile unsigned int*>(&me->max_vals[pass & 1u][lane][0]);
      mh = *reinterpret_cast<const volatile unsigned int*>(&me->max_vals[pass & 1u][lane][1]);
    }
  }
  for (int o = 16; o > 0; o >>= 1) {      // non-negative float bit patterns order like unsigned ints
    mg = max(mg, __shfl_xor_sync(0xffffffffu, mg, o));
    mh = max(mh, __shfl_xor_sync(0xffffffffu, mh, o));
  }
  if (lane == 0) { ctrl->max_abs_g_bits = mg; ctrl->max_abs_h_bits = mh; }
}
// raises this rank's abort word in every rank's mailbox (host error path)
__global__ void k_peer_abort(PeerTable pt) {
  if (threadIdx.x < pt.world) StoreReleaseSys(&pt.mail[threadIdx.x]->abort[pt.rank], 1u);
}

// ------------------------------------------------------------------------------------------------ init
__global__ void k_bt_init(NodeDev* nodes, int* avail, BatchCtrl* ctrl, TreeArraysDev T, int L, int n_local, int n_global) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < L) {
    T.leaf_parent[i] = -1; T.leaf_depth[i] = 0; T.leaf_value[i] = 0.0; T.leaf_weight[i] = 0.0; T.leaf_count[i] = 0;
    avail[i] = -1;
    if (i < L - 1) {
      T.left_child[i] = 0; T.right_child[i] = 0; T.split_feature_inner[i] = 0; T.threshold_in_bin[i] = 0;
      T.internal_count[i] = 0; T.decision_type[i] = 0; T.split_gain[i] = 0.0f; T.internal_value[i] = 0.0; T.internal_weight[i] = 0.0;
    }
  }
  if (i == 0) {
    NodeDev z;
    z.begin = 0; z.count = n_local; z.buf = -1; z.global_count = n_global; z.int_g = z.int_h = 0; z.sum_g = z.sum_h = 0.0;
    z.depth = 0; z.hist_slot = 0; z.state = 0; z.left = z.right = -1; z.pad_ = 0;
    z.cmin = -INFINITY; z.cmax = INFINITY;
    z.child_feat[0] = z.child_feat[1] = -1; z.child_count[0] = z.child_count[1] = 0; z.child_gain[0] = z.child_gain[1] = KMINSCORE;
    z.best = EmptySplit(-1, -1);
    nodes[0] = z;
    avail[0] = 0;
    ctrl->num_leaves = 1; ctrl->done = 0; ctrl->num_nodes = 1; ctrl->num_slots = 1; ctrl->batch_n = 0; ctrl->spec_used = 0;
    ctrl->rounds = 0; ctrl->applied = 0; ctrl->node_draws = 0; ctrl->last_searched = 0;
    for (int k = 0; k < kMaxBatch; ++k) { ctrl->find_count[k][0] = 0; ctrl->find_count[k][1] = 0; ctrl->part_cursor[k] = 0ULL; }
  }
}

// ------------------------------------------------------------------------------------------------ K4 wrapper
// grid = (row CTAs, feature tiles).  ROOT: the CTAs share the rows of node 0.  Otherwise the smaller children of the round are
// cut into chunks of rows_per_cta rows, sized so that there are at most as many chunks as row CTAs; CTA x takes chunk x.  (One
// CTA fits per SM — 160 KB of shared memory — so a grid with a CTA per (chunk, batch slot) would run in several waves of
// mostly idle CTAs; that was 15 us of fixed cost per round.)
// The body: all kHistThreads threads of a CTA; (cta_x, tile) = the CTA's place in a (row_ctas x tiles) arrangement.  The shared
// histograms must be zero on entry and are zero again on return.
struct HistShared { int begin[kMaxBatch], count[kMaxBatch], buf[kMaxBatch]; int chunks[kMaxBatch + 1]; int rpc; unsigned long long t_flush; };

template <int R, int BPL, bool FLT, bool ROOT>
__device__ __forceinline__ void HistBody(const uint8_t* __restrict__ rm, unsigned rm_stride, const uint16_t* __restrict__ gh16,
                                         const int2* __restrict__ ghfix, const int* idx0, const int* idx1, const NodeDev* nodes,
                                         BatchCtrl* ctrl, long long* work, int work_elems, int F, int rows_per_flush, int row_ctas,
                                         int cta_x, int tile, unsigned int* hsm, unsigned lane_addr, HistShared* sh,
                                         const LearnerParams& P) {
  constexpr int LPR = 32 / R;
  constexpr int FT = LPR * BPL;
  const int lane = threadIdx.x & 31;
  const int q = lane % LPR;
  const bool kmask = (tile * FT + q * BPL) < F;
  const uint8_t* rmc = rm + static_cast<size_t>(tile) * FT + (kmask ? q : 0) * BPL;
  asm volatile("" : "+l"(rmc));
  if (threadIdx.x == 0) {
    const int bn = ROOT ? 1 : ctrl->batch_n;
    long long total = 0;
    int active = 0;
    for (int k = 0; k < kMaxBatch; ++k) {
      int cnt = 0;
      if (ROOT) {
        if (k == 0) { sh->begin[0] = 0; sh->buf[0] = -1; cnt = nodes[0].count; active = 1; }
      } else if (k < bn) {
        const BatchItem& it = ctrl->batch[k];
        const ChildGeom g = MakeGeom(it, ctrl->part_cursor[k], P);
        if (g.find) {
          sh->begin[k] = g.left_smaller ? it.src_begin : g.rbegin;
          sh->buf[k] = g.buf;
          cnt = g.left_smaller ? g.nl : g.nr;
          ++active;
        }
      }
      sh->count[k] = cnt;       // an item with an empty smaller child (possible on one GPU of several) needs no CTA
      total += cnt;
    }
    const int avail_ctas = max(1, row_ctas - active);
    long long rpc = (total + avail_ctas - 1) / avail_ctas;
    int rows_per_cta = static_cast<int>(rpc > 0x3fffffff ? 0x3fffffff : rpc);
    rows_per_cta = max((rows_per_cta + 31) & ~31, kHistMinRowsPerCta);
    sh->rpc = rows_per_cta;
    int c = 0;
    for (int k = 0; k < kMaxBatch; ++k) {
      sh->chunks[k] = c;
      const int cnt = sh->count[k];
      c += (cnt + rows_per_cta - 1) / rows_per_cta;
      if (cnt > 0 && cta_x == 0 && tile == 0) {
        atomicAdd(reinterpret_cast<unsigned long long*>(&ctrl->hist_rows), static_cast<unsigned long long>(cnt));
        atomicAdd(reinterpret_cast<unsigned long long*>(&ctrl->hist_launches), 1ULL);
      }
    }
    sh->chunks[kMaxBatch] = c;
  }
  __syncthreads();
  LGB_TRACE(ctrl, 23);      // set-up done
  const bool tracing = c_trace_on != 0 && blockIdx.x == 0 && blockIdx.y == 0;
  if (tracing && threadIdx.x == 0) sh->t_flush = 0ULL;
  const int total_chunks = sh->chunks[kMaxBatch];
  const int rows_per_cta = sh->rpc;
  for (int chunk = cta_x; chunk < total_chunks; chunk += row_ctas) {      // normally one chunk per CTA
    int k = 0;
    while (k + 1 < kMaxBatch && chunk >= sh->chunks[k + 1]) ++k;
    const int begin = sh->begin[k], count = sh->count[k], buf = sh->buf[k];
    const int lo = (chunk - sh->chunks[k]) * rows_per_cta;
    const int hi = min(count, lo + rows_per_cta);
    const int* idx = ROOT ? nullptr : (buf == 0 ? idx0 : idx1);
    HistRows<R, BPL, FLT, ROOT, !FLT>(rmc, rm_stride, gh16, ghfix, idx, begin, lo, hi, lane_addr, hsm,
                                      work + static_cast<size_t>(ROOT ? 0 : k) * work_elems, F, tile, rows_per_flush, kmask,
                                      tracing ? &sh->t_flush : nullptr);
  }
  __syncthreads();      // sh is rewritten by the next call
  if (tracing && threadIdx.x == 0 && ctrl->trace != nullptr && sh->t_flush != 0ULL) {      // id 24: when CTA 0's (last) flush began
    const int i = atomicAdd(&ctrl->trace_n, 1);
    if (i < ctrl->trace_cap) { ctrl->trace[2 * i] = 24; ctrl->trace[2 * i + 1] = static_cast<long long>(sh->t_flush); }
  }
}

template <int R, int BPL, bool FLT, bool ROOT>
__global__ void __launch_bounds__(kHistThreads, 1)
k_bt_hist(const uint8_t* __restrict__ rm, unsigned rm_stride, const uint16_t* __restrict__ gh16, const int2* __restrict__ ghfix,
          const int* __restrict__ idx0, const int* __restrict__ idx1, const NodeDev* __restrict__ nodes,
          BatchCtrl* __restrict__ ctrl, long long* __restrict__ work, int work_elems, int F, int rows_per_flush, int row_ctas,
          LearnerParams P, PeerTable pt) {
  extern __shared__ unsigned int hsm_raw[];
  __shared__ HistShared sh;
  PdlLaunchDependents();
  const unsigned raw_addr = static_cast<unsigned>(__cvta_generic_to_shared(hsm_raw));
  const unsigned sbase = (raw_addr + kHistArrayBytes - 1) & ~(kHistArrayBytes - 1);
  unsigned int* hsm = hsm_raw + ((sbase - raw_addr) >> 2);
  const unsigned lane_addr = sbase + (threadIdx.x & 31) * 4;
  for (int i = threadIdx.x; i < 4 * kBins * 32; i += blockDim.x) hsm[i] = 0u;      // before the wait: overlaps the predecessor
  LGB_TRACE(ctrl, 20);
  PdlWait();
  LGB_TRACE(ctrl, 21);
  if (Finished(ctrl)) return;
  __syncthreads();
  HistBody<R, BPL, FLT, ROOT>(rm, rm_stride, gh16, ghfix, idx0, idx1, nodes, ctrl, work, work_elems, F, rows_per_flush, row_ctas,
                              blockIdx.x, blockIdx.y, hsm, lane_addr, &sh, P);
  if (pt.world > 0) {      // the CTA that finishes last tells the other ranks that this rank's histograms are complete
    __shared__ int s_is_last;
    __threadfence();       // this CTA's global atomics are performed (device scope: the local L2 also serves the peers' reads) ...
    __syncthreads();
    PeerMail* me = pt.mail[pt.rank];
    if (threadIdx.x == 0) {
      const bool last = atomicAdd(&me->hist_count, 1u) == gridDim.x * gridDim.y - 1;
      if (last) me->hist_count = 0u;
      s_is_last = last;
    }
    __syncthreads();
    if (s_is_last && threadIdx.x < pt.world) {
      __threadfence_system();      // ... and the last CTA publishes all of them system-wide before the flag
      StoreReleaseSys(&pt.mail[threadIdx.x]->hist_flag[pt.rank], pt.seq);      // one remote store per peer (own mailbox included)
    }
  }
  LGB_TRACE(ctrl, 22);
}

// ------------------------------------------------------------------------------------------------ K5 + K6
// grid = (ceil(features / 8), root ? 1 : 2, batch slots), 256 threads: WARP w of CTA (x, c, k) is the unit (feature 8 x + w, child c,
// batch item k); c = 0: the freshly built smaller child, c = 1: the larger child = parent - smaller.  A unit loads its feature's
// bins (summed over the ranks when histograms are exchanged over peer memory), stores the child's histogram in its pool slot and
// runs the warp-scan split search (WarpFindBest, kernels.cuh).  With few features (LearnerParams::reduce_in_select) the replay
// kernel picks each child's best feature; otherwise the unit that finishes last for a child reduces over the (rank's) features.
constexpr int kFindWarps = 8;

template <bool QUANT>
__device__ __forceinline__ void FindUnitWarp(const long long* work, int work_elems, long long* pool, const FeatureMetaDev* __restrict__ fmeta,
                                             NodeDev* nodes, BatchCtrl* ctrl, SplitDev* feat_best, unsigned char* splittable,
                                             const LearnerParams& P, int root, int f, int c, int k, const PeerTable& pt) {
  const int F = P.num_features;
  const int lane = threadIdx.x & 31;
  // feature-sliced exchange: this rank owns features [f_first, f_first + nf) and sends the best of them to every rank
  const bool sliced = pt.world > 0 && pt.sliced != 0;
  const int f_first = sliced ? pt.f_lo : 0;
  const int nf = sliced ? pt.f_hi - pt.f_lo : F;
  // ---- which node is this, and what do we know about it
  NodeDev me;              // what gets published for the child (root: only the fields the search needs)
  int node, parent = -1, parent_slot = 0, side = 0;
  bool find = true;
  if (root) {
    node = 0;
    me = nodes[0];
  } else {
    const BatchItem& it = ctrl->batch[k];
    const ChildGeom g = MakeGeom(it, ctrl->part_cursor[k], P);
    const bool is_left = (c == 0) == g.left_smaller;      // c == 0: the smaller child (its histogram was just built)
    node = is_left ? it.left : it.right;
    side = is_left ? 0 : 1;
    parent = it.node;
    parent_slot = it.parent_slot;
    find = g.find;
    me.begin = is_left ? it.src_begin : g.rbegin;
    me.count = is_left ? g.nl : g.nr;
    me.buf = g.buf;
    me.global_count = is_left ? g.lc : g.rc;
    me.int_g = is_left ? it.left_int_g : it.right_int_g; me.int_h = is_left ? it.left_int_h : it.right_int_h;
    me.sum_g = is_left ? it.left_sum_g : it.right_sum_g; me.sum_h = is_left ? it.left_sum_h : it.right_sum_h;
    me.depth = it.child_depth;
    me.cmin = it.cmin; me.cmax = it.cmax;
    if (!it.is_cat && it.mono != 0) {      // BasicLeafConstraints::Update: both ranges are cut at the mean of the two outputs
      const double mid = (it.left_output + it.right_output) / 2.0f;
      if ((it.mono < 0) == is_left) me.cmin = fmax(mid, me.cmin); else me.cmax = fmin(mid, me.cmax);
    }
    me.hist_slot = c == 0 ? it.fresh_slot : it.parent_slot;     // the larger child takes over the parent's slot
    me.state = 0; me.left = me.right = -1; me.pad_ = 0;
  }
  me.child_feat[0] = me.child_feat[1] = -1; me.child_count[0] = me.child_count[1] = 0; me.child_gain[0] = me.child_gain[1] = KMINSCORE;
  if (!find) {         // children that are not searched are published by their first unit alone
    if (f != f_first) return;
    if (lane == 0) {
      me.best = EmptySplit(-1, -1);
      nodes[node] = me;
      nodes[parent].child_gain[side] = KMINSCORE; nodes[parent].child_feat[side] = -1; nodes[parent].child_count[side] = me.global_count;
      __threadfence();
    }
    return;
  }
  const FeatureMetaDev fm = fmeta[f];
  const double gscale = ctrl->grad_scale, hscale = ctrl->hess_scale;
  const size_t slot_stride = static_cast<size_t>(F) * kBins * 2;
  const size_t item_off = static_cast<size_t>(root ? 0 : k) * work_elems;
  const size_t bin0 = static_cast<size_t>(f) * kBins + lane * kBinsPerLane;      // the lane's first bin of this feature
  long long g_in[kBinsPerLane], h_in[kBinsPerLane];
  if (pt.world > 0) {      // the sum of all ranks' bins, read over peer memory
    #pragma unroll
    for (int j = 0; j < kBinsPerLane; ++j) { g_in[j] = 0; h_in[j] = 0; }
    for (int r = 0; r < pt.world; ++r) {
      const long long* w = pt.work[r] + pt.parity_off + item_off;
      if (QUANT) {
        const longlong2* src = reinterpret_cast<const longlong2*>(w + bin0);
        #pragma unroll
        for (int j = 0; j < kBinsPerLane / 2; ++j) { const longlong2 t = src[j]; g_in[2 * j] += t.x; g_in[2 * j + 1] += t.y; }
      } else {
        const longlong2* src = reinterpret_cast<const longlong2*>(w + bin0 * 2);
        #pragma unroll
        for (int j = 0; j < kBinsPerLane; ++j) { const longlong2 t = src[j]; g_in[j] += t.x; h_in[j] += t.y; }
      }
    }
    if (QUANT) {
      #pragma unroll
      for (int j = 0; j < kBinsPerLane; ++j) { const long long a = g_in[j]; h_in[j] = a & 0xffffffffLL; g_in[j] = (a - h_in[j]) >> 32; }
    }
  } else if (QUANT) {     // packed work area: (sum g << 32) + sum h
    const longlong2* src = reinterpret_cast<const longlong2*>(work + item_off + bin0);
    #pragma unroll
    for (int j = 0; j < kBinsPerLane / 2; ++j) {
      const longlong2 t = src[j];
      h_in[2 * j] = t.x & 0xffffffffLL; g_in[2 * j] = (t.x - h_in[2 * j]) >> 32;
      h_in[2 * j + 1] = t.y & 0xffffffffLL; g_in[2 * j + 1] = (t.y - h_in[2 * j + 1]) >> 32;
    }
  } else {
    const longlong2* src = reinterpret_cast<const longlong2*>(work + item_off + bin0 * 2);
    #pragma unroll
    for (int j = 0; j < kBinsPerLane; ++j) { const longlong2 t = src[j]; g_in[j] = t.x; h_in[j] = t.y; }
  }
  bool used = true;
  if (!root) {
    used = splittable[static_cast<size_t>(parent) * F + f] != 0;
    if (c == 1) {
      const longlong2* p = reinterpret_cast<const longlong2*>(pool + static_cast<size_t>(parent_slot) * slot_stride + bin0 * 2);
      #pragma unroll
      for (int j = 0; j < kBinsPerLane; ++j) { const longlong2 t = p[j]; g_in[j] = t.x - g_in[j]; h_in[j] = t.y - h_in[j]; }
    }
  }
  {
    longlong2* dst = reinterpret_cast<longlong2*>(pool + static_cast<size_t>(me.hist_slot) * slot_stride + bin0 * 2);
    #pragma unroll
    for (int j = 0; j < kBinsPerLane; ++j) dst[j] = make_longlong2(g_in[j], h_in[j]);
  }
  SplitDev* fb = feat_best + static_cast<size_t>(root ? 0 : (k * 2 + c)) * F;
  unsigned char* spl = splittable + static_cast<size_t>(node) * F + f;
  long long tg = me.int_g, th = me.int_h;
  // column sample of this tree (feature_fraction): a feature outside it has no split.  At the root it still runs the search,
  // because the leaf totals come out of it, and its result is discarded.
  const bool sampled = P.feature_mask == nullptr || P.feature_mask[f] != 0;
  if (!used || (!sampled && !root)) {
    if (lane == 0) { fb[f] = EmptySplit(f, fm.real_index); *spl = 0; }
  } else {
    WarpFindBest<QUANT>(g_in, h_in, fm, f, root != 0, me.int_g, me.int_h, me.sum_g, me.sum_h, me.global_count, P, gscale, hscale,
                        fb + f, spl, &tg, &th);
    if (!sampled && lane == 0) { fb[f] = EmptySplit(f, fm.real_index); *spl = 0; }      // lane 0 wrote the result it now replaces
  }
  if (P.reduce_in_select && !sliced) {
    // the replay kernel picks the child's best feature (SelectReduceChildren); the first unit publishes everything else
    if (f == 0 && lane == 0) {
      if (root) {        // the root's totals come out of the search itself (every feature sees the same sums)
        me.int_g = tg; me.int_h = th;
        me.sum_g = static_cast<double>(tg) * gscale; me.sum_h = static_cast<double>(th) * hscale;
      }
      NodeDev& N = nodes[node];
      N.begin = me.begin; N.count = me.count; N.buf = me.buf; N.global_count = me.global_count;
      N.int_g = me.int_g; N.int_h = me.int_h; N.sum_g = me.sum_g; N.sum_h = me.sum_h;
      N.depth = me.depth; N.hist_slot = me.hist_slot; N.state = me.state; N.left = me.left; N.right = me.right; N.pad_ = 0;
      N.cmin = me.cmin; N.cmax = me.cmax;
      N.child_feat[0] = N.child_feat[1] = -1; N.child_count[0] = N.child_count[1] = 0;
      N.child_gain[0] = N.child_gain[1] = KMINSCORE;
    }
    return;
  }
  // the unit that finishes last for this child reduces over the (rank's) features and publishes the child
  int last = 0;
  if (lane == 0) {
    __threadfence();
    last = atomicAdd(&ctrl->find_count[root ? 0 : k][c], 1) == nf - 1;
  }
  last = __shfl_sync(0xffffffffu, last, 0);
  if (!last) return;
  __threadfence();
  double bg = KMINSCORE; int bf = -1, bi = -1;
  for (int j = f_first + lane; j < f_first + nf; j += 32) {
    const double g = fb[j].gain; const int rf = fb[j].real_feature;
    if (bi < 0 || SplitBetter(g, rf, bg, bf)) { bg = g; bf = rf; bi = j; }
  }
  for (int o = 16; o > 0; o >>= 1) {
    const double g = __shfl_xor_sync(0xffffffffu, bg, o);
    const int rf = __shfl_xor_sync(0xffffffffu, bf, o);
    const int i = __shfl_xor_sync(0xffffffffu, bi, o);
    if (i >= 0 && (bi < 0 || SplitBetter(g, rf, bg, bf))) { bg = g; bf = rf; bi = i; }
  }
  if (lane == 0) {
    if (root) {        // the root's totals come out of the search itself (every feature sees the same sums)
      me.int_g = tg; me.int_h = th;
      me.sum_g = static_cast<double>(tg) * gscale; me.sum_h = static_cast<double>(th) * hscale;
    }
    if (sliced) {      // the child's best split is chosen by the replay kernel from all ranks' records (SelectReducePeers)
      me.best = EmptySplit(-1, -1);
      nodes[node] = me;
    } else {
      me.best = fb[bi];
      nodes[node] = me;
      if (!root) {
        nodes[parent].child_gain[side] = me.best.gain; nodes[parent].child_feat[side] = me.best.real_feature;
        nodes[parent].child_count[side] = me.global_count;
      }
    }
    ctrl->find_count[root ? 0 : k][c] = 0;
  }
  if (sliced && lane < pt.world) {        // lane r stores this rank's record into rank r's mailbox, then stamps it
    SplitMail* dst = &pt.mail[lane]->split[pt.seq & 1u][root ? 0 : (k * 2 + c)][pt.rank];
    dst->s = fb[bi];
    dst->tot_g = tg; dst->tot_h = th;
    __threadfence_system();
    StoreReleaseSys(&dst->seq, pt.seq);
  }
}

template <bool QUANT>
__global__ void __launch_bounds__(kFindWarps * 32, 2) k_bt_find_warp(const long long* __restrict__ work, int work_elems, long long* __restrict__ pool,
                                                             const FeatureMetaDev* __restrict__ fmeta, NodeDev* __restrict__ nodes,
                                                             BatchCtrl* __restrict__ ctrl, SplitDev* __restrict__ feat_best,
                                                             unsigned char* __restrict__ splittable, LearnerParams P, int root, PeerTable pt) {
  PdlLaunchDependents();
  LGB_TRACE(ctrl, 30);
  PdlWait();
  LGB_TRACE(ctrl, 31);
  if (Finished(ctrl)) return;
  const int k = blockIdx.z;
  if (!root && k >= ctrl->batch_n) return;
  int f = blockIdx.x * kFindWarps + (threadIdx.x >> 5);
  if (pt.world > 0) {      // every rank's histograms of this exchange must be complete: their words arrive in this rank's own mailbox
    __shared__ int s_ok;
    if (threadIdx.x == 0) s_ok = 1;
    __syncthreads();
    if (threadIdx.x < pt.world && !PeerWait(&pt.mail[pt.rank]->hist_flag[threadIdx.x], pt.seq, pt, ctrl)) s_ok = 0;
    __syncthreads();
    if (!s_ok) return;
    if (pt.sliced) f += pt.f_lo;
  }
  const int f_end = (pt.world > 0 && pt.sliced) ? pt.f_hi : P.num_features;
  if (f < f_end) FindUnitWarp<QUANT>(work, work_elems, pool, fmeta, nodes, ctrl, feat_best, splittable, P, root, f, blockIdx.y, k, pt);
  LGB_TRACE(ctrl, 32);
}

// EXT = false: the dataset has neither categorical nor wide features and no column is constrained — that code is compiled out, so
// the plain numerical search is the kernel it was before those features existed (with them merely present the unit was 20 % slower).
// ---- the same search with a 256-thread CTA per unit (one bin per thread, block-wide scans): one threshold per thread makes a
// unit's latency half that of the warp form (4.4 vs 9.4 us, B200), so it is what narrow data uses, where a round has only a few
// hundred units (C2: 28 x 2 x 8) and is bound by latency; the warp form wins when the units no longer fit one wave (C5: 32000).
template <bool QUANT, bool EXT>
__device__ __forceinline__ void FindUnit(const long long* work, int work_elems, long long* pool, const FeatureMetaDev* __restrict__ fmeta,
                                         NodeDev* nodes, BatchCtrl* ctrl, SplitDev* feat_best, unsigned char* splittable,
                                         const LearnerParams& P, int root, int f, int c, int k, int sub, const PeerTable* pt = nullptr) {
  __shared__ int s_last_all[kMaxFindSubs];
  __shared__ double s_gain_all[kMaxFindSubs][8];
  __shared__ int s_feat_all[kMaxFindSubs][8], s_idx_all[kMaxFindSubs][8];
  int& s_last = s_last_all[sub];
  double* s_gain = s_gain_all[sub]; int* s_feat = s_feat_all[sub]; int* s_idx = s_idx_all[sub];
  const int F = P.num_features;
  const int b = threadIdx.x & (kFindThreads - 1);
  // feature-sliced exchange: this rank owns features [f_first, f_first + nf) and sends the best of them to every rank
  const bool sliced = pt != nullptr && pt->world > 0 && pt->sliced != 0;
  const int f_first = sliced ? pt->f_lo : 0;
  const int nf = sliced ? pt->f_hi - pt->f_lo : F;
  // ---- which node is this, and what do we know about it
  NodeDev me;              // what gets published for the child (root: only the fields the search needs)
  int node, parent = -1, parent_slot = 0, side = 0;
  bool find = true;
  if (root) {
    node = 0;
    me = nodes[0];
  } else {
    const BatchItem& it = ctrl->batch[k];
    const ChildGeom g = MakeGeom(it, ctrl->part_cursor[k], P);
    const bool is_left = (c == 0) == g.left_smaller;      // c == 0: the smaller child (its histogram was just built)
    node = is_left ? it.left : it.right;
    side = is_left ? 0 : 1;
    parent = it.node;
    parent_slot = it.parent_slot;
    find = g.find;
    me.begin = is_left ? it.src_begin : g.rbegin;
    me.count = is_left ? g.nl : g.nr;
    me.buf = g.buf;
    me.global_count = is_left ? g.lc : g.rc;
    me.int_g = is_left ? it.left_int_g : it.right_int_g; me.int_h = is_left ? it.left_int_h : it.right_int_h;
    me.sum_g = is_left ? it.left_sum_g : it.right_sum_g; me.sum_h = is_left ? it.left_sum_h : it.right_sum_h;
    me.depth = it.child_depth;
    me.cmin = it.cmin; me.cmax = it.cmax;
    if (!it.is_cat && it.mono != 0) {      // BasicLeafConstraints::Update: both ranges are cut at the mean of the two outputs
      const double mid = (it.left_output + it.right_output) / 2.0f;
      if ((it.mono < 0) == is_left) me.cmin = fmax(mid, me.cmin); else me.cmax = fmin(mid, me.cmax);
    }
    me.hist_slot = c == 0 ? it.fresh_slot : it.parent_slot;     // the larger child takes over the parent's slot
    me.state = 0; me.left = me.right = -1; me.pad_ = 0;
  }
  me.child_feat[0] = me.child_feat[1] = -1; me.child_count[0] = me.child_count[1] = 0; me.child_gain[0] = me.child_gain[1] = KMINSCORE;
  if (!find) {         // children that are not searched are published by their first unit alone
    if (f != f_first) return;
    if (b == 0) {
      me.best = EmptySplit(-1, -1);
      nodes[node] = me;
      nodes[parent].child_gain[side] = KMINSCORE; nodes[parent].child_feat[side] = -1; nodes[parent].child_count[side] = me.global_count;
      __threadfence();
    }
    return;
  }
  const FeatureMetaDev fm = fmeta[f];
  const double gscale = ctrl->grad_scale, hscale = ctrl->hess_scale;
  const size_t slot_stride = static_cast<size_t>(F) * kBins * 2;
  const size_t item_off = static_cast<size_t>(root ? 0 : k) * work_elems;
  // K5 for bin b of column `col`: this rank's (or all ranks') fresh sums, turned into the larger child's by subtracting them from
  // the parent's, and stored in the child's pool slot
  // (explicit value captures: capturing the node record `me` by reference would pin that whole struct in local memory)
  const int my_slot = me.hist_slot;
  const bool subtract = !root && c == 1;
  auto child_bin = [b, item_off, slot_stride, my_slot, subtract, parent_slot, work, pool, pt](int col, bool store) -> longlong2 {
    const size_t off = (static_cast<size_t>(col) * kBins + b) * 2;
    longlong2 v;
    if (pt != nullptr && pt->world > 0) {      // sum of all ranks' bins, read over peer memory
      long long a = 0, bsum = 0;
      for (int r = 0; r < pt->world; ++r) {
        const long long* w = pt->work[r] + pt->parity_off + item_off;
        if (QUANT) a += w[static_cast<size_t>(col) * kBins + b];
        else { const longlong2 t = *reinterpret_cast<const longlong2*>(w + off); a += t.x; bsum += t.y; }
      }
      if (QUANT) { v.y = a & 0xffffffffLL; v.x = (a - v.y) >> 32; }
      else { v.x = a; v.y = bsum; }
    } else if (QUANT) {     // packed work area: (sum g << 32) + sum h
      const long long pk = work[item_off + static_cast<size_t>(col) * kBins + b];
      v.y = pk & 0xffffffffLL;
      v.x = (pk - v.y) >> 32;
    } else {
      v = *reinterpret_cast<const longlong2*>(work + item_off + off);
    }
    if (subtract) {
      const longlong2 p = *reinterpret_cast<const longlong2*>(pool + static_cast<size_t>(parent_slot) * slot_stride + off);
      v.x = p.x - v.x; v.y = p.y - v.y;
    }
    if (store) *reinterpret_cast<longlong2*>(pool + static_cast<size_t>(my_slot) * slot_stride + off) = v;
    return v;
  };
  // a wide feature (more than 256 bins: several columns, dataset.h) is handled whole by the unit of its first column; the units
  // of its other columns have nothing to compute and report "no split"
  // (the loads do not wait for the feature's description: one memory round trip, not two, on the unit's critical path)
  const longlong2 v = child_bin(f, false);
  if (!EXT || fm.wide_cols == 0)
    *reinterpret_cast<longlong2*>(pool + static_cast<size_t>(my_slot) * slot_stride + (static_cast<size_t>(f) * kBins + b) * 2) = v;
  const bool used = root || splittable[static_cast<size_t>(parent) * F + f] != 0;
  SplitDev* fb = feat_best + static_cast<size_t>(root ? 0 : (k * 2 + c)) * F;
  unsigned char* spl = splittable + static_cast<size_t>(node) * F + f;
  long long tg = me.int_g, th = me.int_h;
  // column sample of this tree (feature_fraction): a feature outside it has no split.  At the root it still runs the search,
  // because the leaf totals come out of it, and its result is discarded.
  const bool sampled = P.feature_mask == nullptr || P.feature_mask[f] != 0;
  // feature_fraction_bynode (ColSampler::GetByNode): a feature outside the leaf's own sample is searched all the same — its
  // splittable flag is recorded — but cannot be the leaf's split
  bool in_node_sample = true;
  if (EXT && P.node_mask != nullptr) in_node_sample = P.node_mask[static_cast<size_t>(root ? 0 : ctrl->node_draws + c) * F + f] != 0;
  McCtx mc{0, 0, 0.0, 0.0};
  if (EXT && P.monotone != nullptr) { mc.on = 1; mc.mono = fm.bin_type != 0 ? 0 : P.monotone[f]; mc.cmin = me.cmin; mc.cmax = me.cmax; }
  if (EXT && fm.wide_cols < 0) {
    if (root) WideRootTotals(child_bin, fmeta[f + fm.wide_cols], f + fm.wide_cols, &tg, &th, sub);
    if (b == 0) { fb[f] = EmptySplit(f, fm.real_index); *spl = 0; }
  } else if (EXT && fm.wide_cols > 0) {
    // the columns' sums are needed by the children whatever happens to this feature here (K5 runs for every column)
    BlockFindBestWide<QUANT>(child_bin, fm, f, root != 0, me.int_g, me.int_h, me.sum_g, me.sum_h, me.global_count, P, gscale, hscale,
                             fb + f, spl, &tg, &th, pool + static_cast<size_t>(me.hist_slot) * slot_stride, sub, mc, me.depth);
    if ((!used || !sampled) && b == 0) { fb[f] = EmptySplit(f, fm.real_index); *spl = 0; }
  } else if (!used || (!sampled && !root)) {
    if (b == 0) { fb[f] = EmptySplit(f, fm.real_index); *spl = 0; }
  } else {
    if (EXT && fm.bin_type != 0)
      BlockFindBestCat<QUANT>(v.x, v.y, fm, f, root != 0, me.int_g, me.int_h, me.sum_g, me.sum_h, me.global_count, P, gscale, hscale,
                              fb + f, spl, &tg, &th, sub, mc, me.depth);
    else if (EXT && mc.on)
      BlockFindBest<QUANT, true>(v.x, v.y, fm, f, root != 0, me.int_g, me.int_h, me.sum_g, me.sum_h, me.global_count, P, gscale, hscale,
                                 fb + f, spl, &tg, &th, sub, mc, me.depth);
    else
      BlockFindBest<QUANT>(v.x, v.y, fm, f, root != 0, me.int_g, me.int_h, me.sum_g, me.sum_h, me.global_count, P, gscale, hscale,
                           fb + f, spl, &tg, &th, sub);
    if (!sampled && b == 0) { fb[f] = EmptySplit(f, fm.real_index); *spl = 0; }      // b == 0 wrote the result it now replaces
  }
  if (EXT && !in_node_sample && b == 0) fb[f] = EmptySplit(f, fm.real_index);      // b == 0 wrote the result it now replaces
  if (P.reduce_in_select && !sliced) {
    // the replay kernel picks the child's best feature (SelectReduceChildren); the first unit publishes everything else
    if (f == 0 && b == 0) {
      if (root) {        // the root's totals come out of the search itself (every feature sees the same sums)
        me.int_g = tg; me.int_h = th;
        me.sum_g = static_cast<double>(tg) * gscale; me.sum_h = static_cast<double>(th) * hscale;
      }
      NodeDev& N = nodes[node];
      N.begin = me.begin; N.count = me.count; N.buf = me.buf; N.global_count = me.global_count;
      N.int_g = me.int_g; N.int_h = me.int_h; N.sum_g = me.sum_g; N.sum_h = me.sum_h;
      N.depth = me.depth; N.hist_slot = me.hist_slot; N.state = me.state; N.left = me.left; N.right = me.right; N.pad_ = 0;
      N.cmin = me.cmin; N.cmax = me.cmax;
      N.child_feat[0] = N.child_feat[1] = -1; N.child_count[0] = N.child_count[1] = 0;
      N.child_gain[0] = N.child_gain[1] = KMINSCORE;
    }
    SubSync(sub);
    return;
  }
  __threadfence();
  SubSync(sub);
  if (b == 0) s_last = (atomicAdd(&ctrl->find_count[root ? 0 : k][c], 1) == nf - 1);
  SubSync(sub);
  if (s_last) {       // the last unit of this child reduces over the (rank's) features and publishes the child
    __threadfence();
    double bg = KMINSCORE; int bf = -1, bi = -1;
    for (int j = f_first + b; j < f_first + nf; j += kFindThreads) {
      const double g = fb[j].gain; const int rf = fb[j].real_feature;
      if (bi < 0 || SplitBetter(g, rf, bg, bf)) { bg = g; bf = rf; bi = j; }
    }
    const int lane = b & 31, warp = b >> 5;
    for (int o = 16; o > 0; o >>= 1) {
      const double g = __shfl_xor_sync(0xffffffffu, bg, o);
      const int rf = __shfl_xor_sync(0xffffffffu, bf, o);
      const int i = __shfl_xor_sync(0xffffffffu, bi, o);
      if (i >= 0 && (bi < 0 || SplitBetter(g, rf, bg, bf))) { bg = g; bf = rf; bi = i; }
    }
    if (lane == 0) { s_gain[warp] = bg; s_feat[warp] = bf; s_idx[warp] = bi; }
    SubSync(sub);
    if (b == 0) {
      bg = s_gain[0]; bf = s_feat[0]; bi = s_idx[0];
      for (int w = 1; w < 8; ++w)
        if (s_idx[w] >= 0 && (bi < 0 || SplitBetter(s_gain[w], s_feat[w], bg, bf))) { bg = s_gain[w]; bf = s_feat[w]; bi = s_idx[w]; }
      if (root) {        // the root's totals come out of the search itself (every feature sees the same sums)
        me.int_g = tg; me.int_h = th;
        me.sum_g = static_cast<double>(tg) * gscale; me.sum_h = static_cast<double>(th) * hscale;
      }
      if (sliced) {      // the child's best split is chosen by the replay kernel from all ranks' records (SelectReducePeers)
        me.best = EmptySplit(-1, -1);
        nodes[node] = me;
        s_idx[0] = bi;
      } else {
        me.best = fb[bi];
        nodes[node] = me;
        if (!root) {
          nodes[parent].child_gain[side] = me.best.gain; nodes[parent].child_feat[side] = me.best.real_feature;
          nodes[parent].child_count[side] = me.global_count;
        }
      }
      ctrl->find_count[root ? 0 : k][c] = 0;
    }
    if (sliced) {        // thread r stores this rank's record into rank r's mailbox, then stamps it
      SubSync(sub);
      if (b < pt->world) {
        SplitMail* dst = &pt->mail[b]->split[pt->seq & 1u][root ? 0 : (k * 2 + c)][pt->rank];
        dst->s = fb[s_idx[0]];
        dst->tot_g = tg; dst->tot_h = th;
        __threadfence_system();
        StoreReleaseSys(&dst->seq, pt->seq);
      }
    }
  }
  SubSync(sub);       // the sub-block's shared scratch is reused by its next unit
}

template <bool QUANT, bool EXT>
__global__ void __launch_bounds__(256) k_bt_find(const long long* __restrict__ work, int work_elems, long long* __restrict__ pool,
                                                 const FeatureMetaDev* __restrict__ fmeta, NodeDev* __restrict__ nodes,
                                                 BatchCtrl* __restrict__ ctrl, SplitDev* __restrict__ feat_best,
                                                 unsigned char* __restrict__ splittable, LearnerParams P, int root, PeerTable pt) {
  PdlLaunchDependents();
  LGB_TRACE(ctrl, 30);
  PdlWait();
  LGB_TRACE(ctrl, 31);
  if (Finished(ctrl)) return;
  const int k = blockIdx.z;
  if (!root && k >= ctrl->batch_n) return;
  int f = blockIdx.x;
  if (pt.world > 0) {      // every rank's histograms of this exchange must be complete: their words arrive in this rank's own mailbox
    if (pt.sliced) { f += pt.f_lo; if (f >= pt.f_hi) return; }
    __shared__ int s_ok;
    if (threadIdx.x == 0) s_ok = 1;
    __syncthreads();
    if (threadIdx.x < pt.world && !PeerWait(&pt.mail[pt.rank]->hist_flag[threadIdx.x], pt.seq, pt, ctrl)) s_ok = 0;
    __syncthreads();
    if (!s_ok) return;
  }
  FindUnit<QUANT, EXT>(work, work_elems, pool, fmeta, nodes, ctrl, feat_best, splittable, P, root, f, blockIdx.y, k, 0, &pt);
  LGB_TRACE(ctrl, 32);
}

// ------------------------------------------------------------------------------------------------ K7 (batched, one pass)
// All leaves of the batch are cut into tiles of kPartTileRows (1024) rows; the CTAs walk the concatenated tile list, so the work is
// balanced whatever the leaf sizes are.  A tile reads its index entries once, looks up the split feature's bin of each row
// (column-major copy), reserves room for its lefts and rights with ONE 64-bit atomic on the leaf's cursor (left count in the
// high word, right count in the low word) and writes the entries to the other index buffer: lefts grow from the front of
// the parent's range, rights from its back.  Inside a tile the order is kept; tiles land in the order their atomics arrive.
// Every sum taken over a leaf's rows is an exact integer sum (quantized or fixed-point gradients), so the order of rows
// inside a leaf does not change any result.  Nothing is "published" at the end: the consumers of the round (histogram, split
// search) derive the children's row ranges from the final cursor value (MakeGeom).
constexpr int kPartTileWords = 32;                        // 4 words per warp, 8 warps (measured: 2 -> 1.24, 4 -> 1.19, 8 -> 1.21, 16 -> 1.40 ms per iteration)
constexpr int kPartTileRows = kPartTileWords * 32;

// Tile list of the round (one thread): tiles[k] = first tile of batch item k, tiles[kMaxBatch] = total.
__device__ __forceinline__ void PartTileList(const BatchCtrl* ctrl, int bn, int* tiles) {
  int t = 0;
  for (int k = 0; k < kMaxBatch; ++k) {
    tiles[k] = t;
    if (k < bn) t += max(1, (ctrl->batch[k].src_count + kPartTileRows - 1) / kPartTileRows);
  }
  tiles[kMaxBatch] = t;
}

// Tiles vb, vb + nvb, ... by one 256-thread sub-block (named barriers: several sub-blocks can share a CTA).
// A tile is a chain of three dependent memory round trips (entries -> bins -> cursor atomic) before its stores; what hides
// them is the number of tiles in flight per SM, not anything inside one sub-block.
template <int ES>
__device__ __forceinline__ void PartitionTiles(const uint8_t* __restrict__ cm, long long n_pad, const int* idx0r, const int* idx1r,
                                               int* idx0w, int* idx1w, const uint16_t* __restrict__ gh16,
                                               const int2* __restrict__ ghfix, BatchCtrl* ctrl, int bn, const int* s_tiles, int vb,
                                               int nvb, int sub) {
  constexpr int NW = kPartThreads / 32;
  constexpr int WPW = kPartTileWords / NW;      // words per warp
  __shared__ int s_wl_all[kMaxFindSubs][NW], s_wv_all[kMaxFindSubs][NW];
  __shared__ unsigned long long s_base_all[kMaxFindSubs];
  int* s_wl = s_wl_all[sub]; int* s_wv = s_wv_all[sub];
  unsigned long long& s_base = s_base_all[sub];
  const int t = threadIdx.x & (kPartThreads - 1);
  const int lane = t & 31, warp = t >> 5;
  const int total_tiles = s_tiles[kMaxBatch];
  for (int tile = vb; tile < total_tiles; tile += nvb) {
    int k = 0;
    while (k + 1 < bn && tile >= s_tiles[k + 1]) ++k;
    const BatchItem& it = ctrl->batch[k];
    const int n = it.src_count, begin = it.src_begin, buf = it.src_buf;
    const int thr = it.threshold, nan_bin = it.nan_bin, dl = it.default_left, is_cat = it.is_cat;
    const uint8_t* col = cm + static_cast<size_t>(it.feature) * n_pad;
    const int* idx = buf < 0 ? nullptr : (buf == 0 ? idx0r : idx1r);
    const unsigned int* bag = buf < 0 ? ctrl->bag_bits : nullptr;
    int* dst = (buf < 0 ? 0 : 1 - buf) == 0 ? idx0w : idx1w;
    const int p_base = (tile - s_tiles[k]) * kPartTileRows + warp * (WPW * 32) + lane;
    int4 ent[WPW];
    #pragma unroll
    for (int u = 0; u < WPW; ++u) {
      const int p = p_base + u * 32;
      ent[u] = make_int4(-1, 0, 0, 0);
      if (p < n) {
        if (idx) {
          if (ES == 2) { const int2 e = reinterpret_cast<const int2*>(idx)[begin + p]; ent[u].x = e.x; ent[u].y = e.y; }
          else ent[u] = reinterpret_cast<const int4*>(idx)[begin + p];
        } else if (bag == nullptr || ((bag[(begin + p) >> 5] >> ((begin + p) & 31)) & 1u)) {
          // the root: rows are 0..n-1 (those in the bag), gradients come from the per-row arrays
          const int row = begin + p;
          ent[u].x = row;
          if (ES == 2) {
            const unsigned int g = gh16[row];
            ent[u].y = static_cast<int>((static_cast<unsigned>(static_cast<int>(static_cast<signed char>(g >> 8))) << 16) + (g & 0xffu));
          } else { const int2 g = ghfix[row]; ent[u].y = g.x; ent[u].z = g.y; }
        }
      }
    }
    int bins[WPW];
    #pragma unroll
    for (int u = 0; u < WPW; ++u) bins[u] = ent[u].x >= 0 ? col[ent[u].x] : -1;
    unsigned int lmask[WPW], rmask[WPW];
    int wl = 0, wr = 0;
    #pragma unroll
    for (int u = 0; u < WPW; ++u) {
      const bool valid = ent[u].x >= 0;
      const bool left = valid && (is_cat ? ((it.cat_bits[bins[u] >> 5] >> (bins[u] & 31)) & 1u) != 0u
                                         : ((bins[u] == nan_bin) ? (dl != 0) : (bins[u] <= thr)));
      lmask[u] = __ballot_sync(0xffffffffu, left);
      rmask[u] = __ballot_sync(0xffffffffu, valid && !left);
      wl += __popc(lmask[u]); wr += __popc(rmask[u]);
    }
    if (lane == 0) { s_wl[warp] = wl; s_wv[warp] = wr; }
    SubSync(sub);
    int lefts_before = 0, rights_before = 0, tile_l = 0, tile_r = 0;
    #pragma unroll
    for (int w = 0; w < NW; ++w) {
      const int a = s_wl[w], v = s_wv[w];
      if (w < warp) { lefts_before += a; rights_before += v; }
      tile_l += a; tile_r += v;
    }
    if (t == 0)
      s_base = atomicAdd(&ctrl->part_cursor[k], (static_cast<unsigned long long>(tile_l) << 32) | static_cast<unsigned>(tile_r));
    SubSync(sub);
    const unsigned long long base = s_base;
    const int lbase = static_cast<int>(base >> 32), rbase = static_cast<int>(base & 0xffffffffu);
    // lefts of this tile: [begin + lbase, +tile_l); rights: [begin + n - rbase - tile_r, begin + n - rbase)
    int lpos = begin + lbase + lefts_before;
    int rpos = begin + n - rbase - tile_r + rights_before;
    const unsigned lt = (1u << lane) - 1u;
    #pragma unroll
    for (int u = 0; u < WPW; ++u) {
      if (ent[u].x >= 0) {
        const bool left = (lmask[u] >> lane) & 1u;
        const int pos = left ? lpos + __popc(lmask[u] & lt) : rpos + __popc(rmask[u] & lt);
        if (ES == 2) reinterpret_cast<int2*>(dst)[pos] = make_int2(ent[u].x, ent[u].y);
        else reinterpret_cast<int4*>(dst)[pos] = ent[u];
      }
      lpos += __popc(lmask[u]);
      rpos += __popc(rmask[u]);
    }
  }
}

template <int ES>
__global__ void __launch_bounds__(kPartThreads, 6) k_bt_partition(const uint8_t* __restrict__ cm, long long n_pad,
                                                               const int* __restrict__ idx0r, const int* __restrict__ idx1r,
                                                               int* __restrict__ idx0w, int* __restrict__ idx1w,
                                                               const uint16_t* __restrict__ gh16, const int2* __restrict__ ghfix,
                                                               BatchCtrl* __restrict__ ctrl, long long* __restrict__ work,
                                                               int work_elems) {
  PdlLaunchDependents();
  LGB_TRACE(ctrl, 10);
  PdlWait();
  LGB_TRACE(ctrl, 11);
  if (Finished(ctrl)) return;
  const int bn = ctrl->batch_n;
  __shared__ int s_tiles[kMaxBatch + 1];
  // the histograms of this batch accumulate into work[0..bn) with atomics: clear them here
  for (size_t i = blockIdx.x * blockDim.x + threadIdx.x; i < static_cast<size_t>(bn) * work_elems; i += static_cast<size_t>(gridDim.x) * blockDim.x)
    work[i] = 0;
  if (threadIdx.x == 0) PartTileList(ctrl, bn, s_tiles);
  __syncthreads();
  PartitionTiles<ES>(cm, n_pad, idx0r, idx1r, idx0w, idx1w, gh16, ghfix, ctrl, bn, s_tiles, blockIdx.x, gridDim.x, 0);
  LGB_TRACE(ctrl, 12);
}

// ------------------------------------------------------------------------------------------------ replay + next batch
// One CTA.  Global memory is touched in parallel phases only; the order-dependent part of the replay runs on shared copies:
//   P0  every thread loads one available leaf (its candidate split, whether it is expanded, its children's candidates) and the
//       per-leaf tree arrays;
//   P1  warp 0 replays the sequential leaf choice and records the commit list;
//   P2  thread j writes the order-independent tree fields of commit j and stages the rest;
//   P3  thread 0 applies the order-dependent bookkeeping of Tree::Split (parent/child links, depths, leaf values) in commit order;
//   P4  per-leaf arrays are written back and the next batch is chosen.
struct SelectSmem {
  double *gain, *cgain0, *cgain1, *lvalue, *lweight, *p_lo, *p_ro, *p_lw, *p_rw;
  int *feat, *node, *rank, *order, *cfeat0, *cfeat1, *cnode0, *cnode1, *lparent, *ldepth, *lcount, *c_leaf, *c_node, *p_lc, *p_rc, *p_feature,
      *p_threshold, *p_flags;
  float* p_gain;
  unsigned char *exp, *isleft;
};
constexpr size_t kSelectSmemPerLeaf = 9 * sizeof(double) + 18 * sizeof(int) + sizeof(float) + 2;
__host__ __device__ inline size_t SelectSmemBytes(int L) { return kSelectSmemPerLeaf * static_cast<size_t>((L + 7) & ~7) + 64; }

// The body: one 256-thread sub-block, synchronised with its named barrier.
//   P0  every thread loads one available leaf: its candidate split and, if it is expanded, what a commit needs (the split, the
//       children's candidates and counts — all in the leaf's own node record) plus the per-leaf tree arrays;
//   P1  warp 0 replays the sequential leaf choice and records the commit list;
//   P2  thread j writes the order-independent tree fields of commit j (from shared memory);
//   P3  thread 0 applies the order-dependent bookkeeping of Tree::Split (parent/child links, depths, leaf values) in commit order;
//   P4  per-leaf arrays are written back and the next batch is chosen.
// S.rank[i] = how many of the nl current leaves come before leaf i in the sequential learner's order (gain descending, real feature
// ascending, slot ascending): all pairs, shared out over the sub-block's 256 threads.
__device__ __forceinline__ void SelectRanks(const SelectSmem& S, int nl, int tid, int sub) {
  for (int i = tid; i < nl; i += kFindThreads) S.rank[i] = 0;
  SubSync(sub);
  for (int i = tid & 31; i < nl; i += 32) {
    const double g = S.gain[i]; const int rf = S.feat[i];
    int cnt = 0;
    for (int j = tid >> 5; j < nl; j += kFindThreads / 32) {
      const double gj = S.gain[j]; const int fj = S.feat[j];
      cnt += (j != i) && (SplitBetter(gj, fj, g, rf) || (!SplitBetter(g, rf, gj, fj) && j < i));
    }
    if (cnt) atomicAdd(&S.rank[i], cnt);
  }
  SubSync(sub);
}

// LearnerParams::reduce_in_select: one warp per child of the round that was searched — its best feature out of the per-feature
// results the split search left in feat_best (ties: the lower real feature, as the search's own reduction), copied into the
// child's node record, and the child's summary into its parent's.
__device__ __forceinline__ void SelectReduceChildren(NodeDev* nodes, const BatchCtrl* ctrl, const SplitDev* feat_best,
                                                     const LearnerParams& P, int after_root, int tid) {
  const int F = P.num_features;
  const int lane = tid & 31, warp = tid >> 5;
  const int units = after_root ? 1 : 2 * ctrl->batch_n;
  for (int kc = warp; kc < units; kc += kFindThreads / 32) {
    const int k = kc >> 1, c = kc & 1;
    int node = 0, parent = -1, side = 0, count = 0;
    if (!after_root) {
      const BatchItem& it = ctrl->batch[k];
      const ChildGeom g = MakeGeom(it, ctrl->part_cursor[k], P);
      if (!g.find) continue;         // published by the split search itself (no candidates)
      const bool is_left = (c == 0) == g.left_smaller;
      node = is_left ? it.left : it.right;
      side = is_left ? 0 : 1;
      parent = it.node;
      count = is_left ? g.lc : g.rc;
    }
    // lane j holds feature j's whole result (F <= 32), so the winner stores it without a second trip to memory
    const SplitDev* fb = feat_best + static_cast<size_t>(kc) * F;
    SplitDev mine;
    double bg = KMINSCORE; int bf = -1, bi = -1;
    if (lane < F) { mine = fb[lane]; bg = mine.gain; bf = mine.real_feature; bi = lane; }
    for (int o = 16; o > 0; o >>= 1) {
      const double g = __shfl_xor_sync(0xffffffffu, bg, o);
      const int rf = __shfl_xor_sync(0xffffffffu, bf, o);
      const int i = __shfl_xor_sync(0xffffffffu, bi, o);
      if (i >= 0 && (bi < 0 || SplitBetter(g, rf, bg, bf))) { bg = g; bf = rf; bi = i; }
    }
    if (lane == bi) nodes[node].best = mine;
    if (lane == 0 && parent >= 0) {
      nodes[parent].child_gain[side] = bg; nodes[parent].child_feat[side] = bf; nodes[parent].child_count[side] = count;
    }
  }
}

// Feature-sliced exchange: one warp per child of the round that was searched — lane r waits for rank r's record of the child in
// this rank's own mailbox (stored there by rank r's split search) and the warp takes the best of them (ties: the lower real
// feature; the ranks' feature ranges are disjoint).  Every rank sees the same G records, so every rank makes the same choice.
__device__ __forceinline__ void SelectReducePeers(NodeDev* nodes, BatchCtrl* ctrl, const LearnerParams& P, const PeerTable& pt,
                                                  int after_root, int tid) {
  const int lane = tid & 31, warp = tid >> 5;
  const int units = after_root ? 1 : 2 * ctrl->batch_n;
  const PeerMail* me = pt.mail[pt.rank];
  for (int kc = warp; kc < units; kc += kFindThreads / 32) {
    const int k = kc >> 1, c = kc & 1;
    int node = 0, parent = -1, side = 0, count = 0;
    if (!after_root) {
      const BatchItem& it = ctrl->batch[k];
      const ChildGeom g = MakeGeom(it, ctrl->part_cursor[k], P);
      if (!g.find) continue;         // published by the split search itself (no candidates)
      const bool is_left = (c == 0) == g.left_smaller;
      node = is_left ? it.left : it.right;
      side = is_left ? 0 : 1;
      parent = it.node;
      count = is_left ? g.lc : g.rc;
    }
    SplitDev mine;
    double bg = KMINSCORE; int bf = -1, bi = -1;
    if (lane < pt.world) {
      const SplitMail* rec = &me->split[pt.seq & 1u][kc][lane];
      if (PeerWait(&rec->seq, pt.seq, pt, ctrl)) { mine = rec->s; bg = mine.gain; bf = mine.real_feature; bi = lane; }
    }
    for (int o = 16; o > 0; o >>= 1) {
      const double g = __shfl_xor_sync(0xffffffffu, bg, o);
      const int rf = __shfl_xor_sync(0xffffffffu, bf, o);
      const int i = __shfl_xor_sync(0xffffffffu, bi, o);
      if (i >= 0 && (bi < 0 || SplitBetter(g, rf, bg, bf) || (!SplitBetter(bg, bf, g, rf) && i < bi))) { bg = g; bf = rf; bi = i; }
    }
    if (bi >= 0 && lane == bi) nodes[node].best = mine;
    if (lane == 0 && parent >= 0) {
      nodes[parent].child_gain[side] = bg; nodes[parent].child_feat[side] = bf; nodes[parent].child_count[side] = count;
    }
  }
}

__device__ __forceinline__ void SelectBody(NodeDev* nodes, int* avail, BatchCtrl* ctrl, const TreeArraysDev& T,
                                           const FeatureMetaDev* __restrict__ fmeta, const LearnerParams& P, int max_nodes,
                                           int spec_budget, int max_batch, double* sel_smem_raw, int sub,
                                           const SplitDev* feat_best, int after_root, const PeerTable& pt) {
  __shared__ int s_state, s_nl, s_m, s_count;
  const int L = P.num_leaves;
  const int Lc = (L + 7) & ~7;
  SelectSmem S;
  {
    double* d = sel_smem_raw;
    S.gain = d; d += Lc; S.cgain0 = d; d += Lc; S.cgain1 = d; d += Lc; S.lvalue = d; d += Lc; S.lweight = d; d += Lc;
    S.p_lo = d; d += Lc; S.p_ro = d; d += Lc; S.p_lw = d; d += Lc; S.p_rw = d; d += Lc;
    int* q = reinterpret_cast<int*>(d);
    S.feat = q; q += Lc; S.node = q; q += Lc; S.rank = q; q += Lc; S.order = q; q += Lc; S.cfeat0 = q; q += Lc; S.cfeat1 = q; q += Lc;
    S.cnode0 = q; q += Lc; S.cnode1 = q; q += Lc; S.lparent = q; q += Lc; S.ldepth = q; q += Lc; S.lcount = q; q += Lc;
    S.c_leaf = q; q += Lc; S.c_node = q; q += Lc; S.p_lc = q; q += Lc; S.p_rc = q; q += Lc; S.p_feature = q; q += Lc;
    S.p_threshold = q; q += Lc; S.p_flags = q; q += Lc;
    S.p_gain = reinterpret_cast<float*>(q); q += Lc;
    unsigned char* c = reinterpret_cast<unsigned char*>(q);
    S.exp = c; c += Lc; S.isleft = c;
  }
  const int tid = threadIdx.x & (kFindThreads - 1);
  if (P.node_mask != nullptr && tid == 0) {      // the searches that came before this replay used up their by-node samples
    if (after_root) { ctrl->node_draws = 1; ctrl->last_searched = 1; }
    else if (ctrl->batch_n > 0) {
      const int searched = MakeGeom(ctrl->batch[0], ctrl->part_cursor[0], P).find ? 1 : 0;
      ctrl->last_searched = searched;
      if (searched) ctrl->node_draws += 2;
    }
  }
  const int nl0 = ctrl->num_leaves;
  // the first leaf of every thread: what does not depend on this round's split search is requested before the reduction below,
  // so those round trips overlap it
  int pre_nd = 0, pre_par = -1, pre_depth = 0, pre_count = 0;
  double pre_value = 0.0, pre_weight = 0.0;
  if (tid < nl0) {
    pre_nd = avail[tid]; pre_par = T.leaf_parent[tid]; pre_depth = T.leaf_depth[tid]; pre_value = T.leaf_value[tid];
    pre_weight = T.leaf_weight[tid]; pre_count = T.leaf_count[tid];
  }
  if (pt.world > 0 && pt.sliced) {
    SelectReducePeers(nodes, ctrl, P, pt, after_root, tid);
    SubSync(sub);
    if (ctrl->abort) return;
  } else if (P.reduce_in_select) {
    SelectReduceChildren(nodes, ctrl, feat_best, P, after_root, tid);
    SubSync(sub);      // P0 reads the node records just completed (same CTA: the barrier orders them)
  }
  // ---- P0
  for (int i = tid; i < nl0; i += kFindThreads) {
    const bool first = i == tid;
    const int nd = first ? pre_nd : avail[i];
    const NodeDev& N = nodes[nd];
    S.node[i] = nd; S.gain[i] = N.best.gain; S.feat[i] = N.best.real_feature;
    const bool e = N.state == 1;
    S.exp[i] = e;
    if (e) {
      S.cnode0[i] = N.left; S.cnode1[i] = N.right;
      S.cgain0[i] = N.child_gain[0]; S.cfeat0[i] = N.child_feat[0];
      S.cgain1[i] = N.child_gain[1]; S.cfeat1[i] = N.child_feat[1];
      const double lo = N.best.left_output, ro = N.best.right_output;
      S.p_lo[i] = (lo != lo) ? 0.0 : lo; S.p_ro[i] = (ro != ro) ? 0.0 : ro;
      S.p_lw[i] = N.best.left_sum_h; S.p_rw[i] = N.best.right_sum_h;
      S.p_lc[i] = N.child_count[0]; S.p_rc[i] = N.child_count[1];
      S.p_feature[i] = N.best.feature; S.p_threshold[i] = N.best.threshold;
      S.p_flags[i] = (N.best.is_cat ? 1 : 0) | (N.best.default_left ? 2 : 0) | ((N.best.missing_type & 3) << 2);      // = Tree's decision_type
      S.p_gain[i] = static_cast<float>(N.best.gain);
    }
    const int par = first ? pre_par : T.leaf_parent[i];
    S.lparent[i] = par; S.ldepth[i] = first ? pre_depth : T.leaf_depth[i]; S.lvalue[i] = first ? pre_value : T.leaf_value[i];
    S.lweight[i] = first ? pre_weight : T.leaf_weight[i];
    S.lcount[i] = first ? pre_count : T.leaf_count[i];
    S.isleft[i] = par >= 0 ? (T.left_child[par] == ~i) : 0;
  }
  SubSync(sub);
  LGB_TRACE(ctrl, 43);
  // ---- P1: replay the sequential order as far as the already expanded leaves allow.  Children of a committed leaf are never
  //      expanded yet, so the replay commits the initial leaves in rank order until the best entry is not an expanded initial
  //      leaf: rank them once (all threads), then one thread walks the ranks against the best child inserted so far.
  SelectRanks(S, nl0, tid, sub);
  for (int i = tid; i < nl0; i += kFindThreads) S.order[S.rank[i]] = i;
  SubSync(sub);
  if (tid == 0) {
    int nl = nl0, m = 0, state = 0;
    double cg = KMINSCORE; int cf = -1, cs = -1;      // best child so far: gain, real feature, slot
    while (true) {
      const int s = m < nl0 ? S.order[m] : -1;       // the m-th best initial leaf (every earlier one was committed)
      bool leaf_best = s >= 0;
      double bg = KMINSCORE;
      if (s >= 0) {
        const double g = S.gain[s]; const int rf = S.feat[s];
        bg = g;
        if (cs >= 0 && (SplitBetter(cg, cf, g, rf) || (!SplitBetter(g, rf, cg, cf) && cs < s))) { leaf_best = false; bg = cg; }
      } else {
        bg = cg;
      }
      if (nl >= L || !(bg > 0.0)) { state = 1; break; }
      if (!leaf_best || !S.exp[s]) { state = 2; break; }
      // commit: the left child keeps the leaf id, the right child becomes leaf `nl`
      S.c_leaf[m] = s; S.c_node[m] = S.node[s];
      const int l = S.cnode0[s], r = S.cnode1[s];
      const double gl = S.cgain0[s], gr = S.cgain1[s];
      const int fl = S.cfeat0[s], fr = S.cfeat1[s];
      S.node[s] = l; S.gain[s] = gl; S.feat[s] = fl; S.exp[s] = 0;
      S.node[nl] = r; S.gain[nl] = gr; S.feat[nl] = fr; S.exp[nl] = 0;
      if (cs < 0 || SplitBetter(gl, fl, cg, cf) || (!SplitBetter(cg, cf, gl, fl) && s < cs)) { cg = gl; cf = fl; cs = s; }
      if (SplitBetter(gr, fr, cg, cf) || (!SplitBetter(cg, cf, gr, fr) && nl < cs)) { cg = gr; cf = fr; cs = nl; }
      ++nl; ++m;
    }
    s_state = state; s_nl = nl; s_m = m;
  }
  SubSync(sub);
  LGB_TRACE(ctrl, 44);
  const int nl = s_nl, m = s_m;
  // ---- P2: order-independent fields of every commit (a leaf id commits at most once per replay, so the commit data loaded
  //      for it in P0 is still the committed node's)
  for (int j = tid; j < m; j += kFindThreads) {
    const int best_leaf = S.c_leaf[j];
    const int new_node = nl0 - 1 + j, new_leaf = nl0 + j;
    T.split_feature_inner[new_node] = S.p_feature[best_leaf];
    T.split_gain[new_node] = S.p_gain[best_leaf];
    T.left_child[new_node] = ~best_leaf;
    T.right_child[new_node] = ~new_leaf;
    T.internal_weight[new_node] = S.p_lw[best_leaf] + S.p_rw[best_leaf];
    T.internal_count[new_node] = S.p_lc[best_leaf] + S.p_rc[best_leaf];
    T.decision_type[new_node] = static_cast<signed char>(S.p_flags[best_leaf]);
    T.threshold_in_bin[new_node] = S.p_threshold[best_leaf];
    if (T.cat_bits != nullptr) {      // a categorical node's bins that go left (all zero for a numerical node)
      const SplitDev& cb = nodes[S.c_node[j]].best;
      #pragma unroll
      for (int w = 0; w < 8; ++w) T.cat_bits[new_node * 8 + w] = cb.is_cat ? cb.cat_bits[w] : 0u;
    }
    nodes[S.c_node[j]].state = 2;
  }
  // ---- P3: order-dependent bookkeeping (reads only what P0 / P1 left in shared memory)
  if (tid == 0) {
    for (int j = 0; j < m; ++j) {
      const int best_leaf = S.c_leaf[j], new_node = nl0 - 1 + j, new_leaf = nl0 + j;
      const int parent = S.lparent[best_leaf];
      if (parent >= 0) {
        if (S.isleft[best_leaf]) T.left_child[parent] = new_node; else T.right_child[parent] = new_node;
      }
      T.internal_value[new_node] = S.lvalue[best_leaf];
      S.lparent[best_leaf] = new_node; S.lparent[new_leaf] = new_node;
      S.isleft[best_leaf] = 1; S.isleft[new_leaf] = 0;
      S.lvalue[new_leaf] = S.p_ro[best_leaf]; S.lweight[new_leaf] = S.p_rw[best_leaf]; S.lcount[new_leaf] = S.p_rc[best_leaf];
      S.lvalue[best_leaf] = S.p_lo[best_leaf]; S.lweight[best_leaf] = S.p_lw[best_leaf]; S.lcount[best_leaf] = S.p_lc[best_leaf];
      S.ldepth[new_leaf] = S.ldepth[best_leaf] + 1;
      S.ldepth[best_leaf] += 1;
    }
  }
  SubSync(sub);
  LGB_TRACE(ctrl, 45);
  // ---- P4: write the per-leaf arrays back
  if (m > 0) {
    for (int i = tid; i < nl; i += kFindThreads) {
      T.leaf_parent[i] = S.lparent[i]; T.leaf_depth[i] = S.ldepth[i]; T.leaf_value[i] = S.lvalue[i];
      T.leaf_weight[i] = S.lweight[i]; T.leaf_count[i] = S.lcount[i];
      avail[i] = S.node[i];
    }
  }
  if (s_state == 1) {
    if (tid == 0) { ctrl->done = 1; ctrl->num_leaves = nl; ctrl->batch_n = 0; }
    return;
  }
  // ---- next batch: the best not-yet-expanded leaves whose rank is below the remaining leaf budget
  const int budget = L - nl;
  SelectRanks(S, nl, tid, sub);
  LGB_TRACE(ctrl, 46);
  // expansions that are not committed yet (the batch's first leaf is certain to be committed by the next replay, the others
  // are speculative): bounding them bounds the nodes / pool slots a tree can consume to (L - 1) + spec_budget expansions
  const int num_nodes = ctrl->num_nodes, num_slots = ctrl->num_slots;
  const int outstanding = (num_nodes - 1) / 2 - (nl - 1);
  const int spec_left = max(0, spec_budget - outstanding);
  const int node_room = (max_nodes - num_nodes) / 2;
  int kmax = min(min(max_batch, 1 + spec_left), node_room);
  if (kmax < 1) kmax = 1;
  // the first kmax eligible leaves in rank order: warp 0 walks the ranks 32 at a time, a ballot gives every eligible leaf its place
  for (int i = tid; i < nl; i += kFindThreads) S.order[S.rank[i]] = i;
  SubSync(sub);
  const int rlim = min(nl, budget);
  int taken = 0;
  for (int r0 = 0; tid < 32 && r0 < rlim && taken < kmax; r0 += 32) {
    const int r = r0 + tid;
    const int i = r < rlim ? S.order[r] : -1;
    const bool eligible = i >= 0 && !S.exp[i] && S.gain[i] > 0.0;
    const unsigned emask = __ballot_sync(0xffffffffu, eligible);
    const int pos = taken + __popc(emask & ((1u << tid) - 1u));
    taken += __popc(emask);
    if (eligible && pos < kmax) {
      const int nd = S.node[i];
      const NodeDev& par = nodes[nd];
      const SplitDev& bs = par.best;
      BatchItem it;
      it.node = nd; it.left = num_nodes + 2 * pos; it.right = it.left + 1;
      it.feature = bs.feature; it.threshold = bs.threshold; it.default_left = bs.default_left;
      it.nan_bin = bs.nan_bin;
      it.is_cat = bs.is_cat; it.pad_ = 0;
      it.mono = bs.monotone_type; it.pad2_ = 0; it.cmin = par.cmin; it.cmax = par.cmax;
      it.left_output = bs.left_output; it.right_output = bs.right_output;
      #pragma unroll
      for (int w = 0; w < 8; ++w) it.cat_bits[w] = bs.is_cat ? bs.cat_bits[w] : 0u;      // (read only for a categorical split)
      it.src_begin = par.begin; it.src_count = par.count; it.src_buf = par.buf;
      it.parent_slot = par.hist_slot; it.fresh_slot = num_slots + pos;
      it.child_depth = par.depth + 1;
      it.search = !(P.max_depth > 0 && it.child_depth >= P.max_depth);
      it.est_left = bs.left_count; it.est_right = bs.right_count;
      it.left_int_g = bs.left_int_g; it.left_int_h = bs.left_int_h; it.right_int_g = bs.right_int_g; it.right_int_h = bs.right_int_h;
      it.left_sum_g = bs.left_sum_g; it.left_sum_h = bs.left_sum_h; it.right_sum_g = bs.right_sum_g; it.right_sum_h = bs.right_sum_h;
      nodes[nd].state = 1; nodes[nd].left = it.left; nodes[nd].right = it.right;      // expanded by the time the next replay runs
      ctrl->batch[pos] = it;
      atomicAdd(reinterpret_cast<unsigned long long*>(&ctrl->part_rows), static_cast<unsigned long long>(it.src_count));
    }
  }
  if (tid == 0) s_count = min(taken, kmax);
  SubSync(sub);
  LGB_TRACE(ctrl, 47);
  if (tid == 0) {
    const int bn = s_count;
    ctrl->batch_n = bn;
    ctrl->num_nodes = num_nodes + 2 * bn;
    ctrl->num_slots = num_slots + bn;
    ctrl->spec_used += bn - 1;      // statistics only
    ctrl->num_leaves = nl;
    ctrl->rounds += 1;
   
    for (int k = 0; k < kMaxBatch; ++k) ctrl->part_cursor[k] = 0ULL;
  }
}

__global__ void __launch_bounds__(256) k_bt_select(NodeDev* __restrict__ nodes, int* __restrict__ avail, BatchCtrl* __restrict__ ctrl,
                                                   TreeArraysDev T, const FeatureMetaDev* __restrict__ fmeta, LearnerParams P,
                                                   int max_nodes, int spec_budget, int max_batch, const SplitDev* feat_best,
                                                   int after_root, PeerTable pt) {
  PdlLaunchDependents();
  LGB_TRACE(ctrl, 40);
  PdlWait();
  LGB_TRACE(ctrl, 41);
  if (Finished(ctrl)) return;
  extern __shared__ double sel_smem_raw[];
  SelectBody(nodes, avail, ctrl, T, fmeta, P, max_nodes, spec_budget, max_batch, sel_smem_raw, 0, feat_best, after_root, pt);
  LGB_TRACE(ctrl, 42);
}

// ------------------------------------------------------------------------------------------------ GOSS
// Gradient-based one-side sampling (upstream src/boosting/goss.hpp, BaggingHelper), one CTA per block of 1024 rows: the top_rate
// share of the block with the largest sum over classes of |g * h| is kept, other_rate of the rest is drawn sequentially from the
// block's LCG stream (the chance of a row depends on how many were drawn before it) and has its gradients scaled by
// (cnt - top_k) / other_k.  Writes the bag as a bit per row and adds the bag size to *count.
__global__ void __launch_bounds__(256) k_goss_sample(float* __restrict__ grad, float* __restrict__ hess, int n, int K, double top_rate,
                                                     double other_rate, unsigned int* __restrict__ rand_state, unsigned int* __restrict__ bits,
                                                     int* __restrict__ count) {
  __shared__ float s_mag[1024];
  __shared__ unsigned int s_bag[32], s_mul[32];
  __shared__ float s_threshold;
  const int blk = blockIdx.x;
  const int start = blk * 1024;
  const int cnt = min(1024, n - start);
  for (int i = threadIdx.x; i < 1024; i += blockDim.x) {
    float a = 0.0f;
    if (i < cnt) for (int k = 0; k < K; ++k) a += fabsf(grad[static_cast<size_t>(k) * n + start + i] * hess[static_cast<size_t>(k) * n + start + i]);
    s_mag[i] = a;
  }
  if (threadIdx.x < 32) { s_bag[threadIdx.x] = 0u; s_mul[threadIdx.x] = 0u; }
  __syncthreads();
  int top_k = static_cast<int>(cnt * top_rate);
  const int other_k = static_cast<int>(cnt * other_rate);
  top_k = max(1, top_k);
  // the top_k-th largest magnitude (ArrayArgs::ArgMaxAtK): the value whose rank, ties broken by position, is top_k - 1
  for (int i = threadIdx.x; i < cnt; i += blockDim.x) {
    const float v = s_mag[i];
    int rank = 0;
    for (int j = 0; j < cnt; ++j) { const float u = s_mag[j]; rank += (u > v) || (u == v && j < i); }
    if (rank == top_k - 1) s_threshold = v;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const float threshold = s_threshold;
    unsigned int x = rand_state[blk];
    int left = 0, big = 0;
    for (int i = 0; i < cnt; ++i) {
      if (s_mag[i] >= threshold) { s_bag[i >> 5] |= 1u << (i & 31); ++left; ++big; continue; }
      const int sampled = left - big;
      const int rest_need = other_k - sampled;
      const int rest_all = (cnt - i) - (top_k - big);
      const double prob = rest_need / static_cast<double>(rest_all);
      x = 214013u * x + 2531011u;
      const float u = static_cast<float>((x >> 16) & 0x7FFF) / 32768.0f;
      if (static_cast<double>(u) < prob) { s_bag[i >> 5] |= 1u << (i & 31); s_mul[i >> 5] |= 1u << (i & 31); ++left; }
    }
    rand_state[blk] = x;
    atomicAdd(count, left);
  }
  __syncthreads();
  const float multiply = static_cast<float>(cnt - top_k) / other_k;
  for (int i = threadIdx.x; i < cnt; i += blockDim.x) {
    if ((s_mul[i >> 5] >> (i & 31)) & 1u) {
      for (int k = 0; k < K; ++k) {
        const size_t idx = static_cast<size_t>(k) * n + start + i;
        grad[idx] *= multiply; hess[idx] *= multiply;
      }
    }
  }
  if (threadIdx.x < 32) bits[blk * 32 + threadIdx.x] = s_bag[threadIdx.x];
}
// every row in the bag (GOSS's first 1 / learning_rate iterations)
__global__ void k_bag_all(unsigned int* __restrict__ bits, int n) {
  const int w = blockIdx.x * blockDim.x + threadIdx.x;
  const int nw = (n + 31) / 32;
  if (w >= ((n + 1023) / 1024) * 32) return;
  unsigned int v = 0u;
  if (w < nw) v = (w == nw - 1 && (n & 31)) ? ((1u << (n & 31)) - 1u) : 0xffffffffu;
  bits[w] = v;
}
// score = (score * pre + v) * post for every row (a random forest's constant tree entering the running mean)
__global__ void k_mix_constant(double* p, long long n, double pre, double v, double post) {
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += static_cast<long long>(gridDim.x) * blockDim.x)
    p[i] = (p[i] * pre + v) * post;
}
__global__ void k_fill_double(double* p, long long n, double v) {
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += static_cast<long long>(gridDim.x) * blockDim.x) p[i] = v;
}

// ------------------------------------------------------------------------------------------------ leaf re-fit (l1 / quantile / mape)
// SerialTreeLearner::RenewTreeOutput with the objective's RenewTreeOutput (upstream regression_objective.hpp): once the tree is
// finished, every leaf's output becomes the (weighted) percentile of label - score over the leaf's rows.  Three steps on the
// device: the residuals (and weights) of the leaves' rows are gathered leaf after leaf (k_renew_collect; the offsets come from the
// node table), sorted per leaf (CUB segmented radix sort, stable), and one warp per leaf evaluates PercentileFun /
// WeightedPercentileFun on its sorted segment (k_renew_pick).
__global__ void k_renew_offsets(const NodeDev* __restrict__ nodes, const int* __restrict__ avail, const BatchCtrl* __restrict__ ctrl, int L,
                                int* __restrict__ off /*[L + 1]*/) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  const int nl = (ctrl->done && !ctrl->abort) ? ctrl->num_leaves : 0;
  int o = 0;
  for (int l = 0; l < L; ++l) {
    off[l] = o;
    if (l < nl && nl > 1) o += nodes[avail[l]].count;
  }
  off[L] = o;
}
template <int ES>
__global__ void k_renew_collect(const NodeDev* __restrict__ nodes, const int* __restrict__ avail, const BatchCtrl* __restrict__ ctrl,
                                const int* __restrict__ off, const int* __restrict__ idx0, const int* __restrict__ idx1,
                                const float* __restrict__ label, const double* __restrict__ score, const float* __restrict__ weight,
                                double* __restrict__ res, float* __restrict__ wout, int use_const_pred, double const_pred) {
  const int l = blockIdx.x;
  if (!ctrl->done || ctrl->abort || l >= ctrl->num_leaves || ctrl->num_leaves <= 1) return;
  const NodeDev& nd = nodes[avail[l]];
  const int* idx = nd.buf == 0 ? idx0 : idx1;
  const int o = off[l];
  for (int p = blockIdx.y * blockDim.x + threadIdx.x; p < nd.count; p += gridDim.y * blockDim.x) {
    const int row = idx[static_cast<size_t>(nd.begin + p) * ES];
    // (a random forest re-fits against label - initial score: RF::TrainOneIter's residual_getter)
    res[o + p] = static_cast<double>(label[row]) - (use_const_pred ? const_pred : score[row]);
    if (weight != nullptr) wout[o + p] = weight[row];
  }
}
// sorted: the leaf's residuals ascending (and their weights).  alpha_is_float: the quantile objective keeps alpha as a float, so
// (1.0f - alpha) * cnt is float arithmetic there; l1 / mape use the double 0.5.
__global__ void k_renew_pick(const BatchCtrl* __restrict__ ctrl, const int* __restrict__ off, const double* __restrict__ sorted,
                             const float* __restrict__ wsorted, double alpha, int alpha_is_float, double* __restrict__ leaf_value,
                             double* __restrict__ out /*[2][L]: value, has rows*/, int L) {
  const int l = blockIdx.x * blockDim.x + threadIdx.x;
  if (l >= L) return;
  out[l] = 0.0; out[L + l] = 0.0;
  if (!ctrl->done || ctrl->abort || l >= ctrl->num_leaves || ctrl->num_leaves <= 1) return;
  const int cnt = off[l + 1] - off[l];
  if (cnt <= 0) { leaf_value[l] = 0.0; return; }      // (data-parallel: another machine holds this leaf's rows)
  const double* d = sorted + off[l];
  double v;
  if (cnt <= 1) {
    v = d[0];
  } else if (wsorted == nullptr) {
    // PercentileFun: with the data DEscending, [pos - 1] and [pos]; here ascending, so [cnt - pos] and [cnt - pos - 1]
    const double float_pos = alpha_is_float ? static_cast<double>((1.0f - static_cast<float>(alpha)) * cnt) : (1.0f - alpha) * cnt;
    const int pos = static_cast<int>(float_pos);
    if (pos < 1) v = d[cnt - 1];
    else if (pos >= cnt) v = d[0];
    else {
      const double bias = float_pos - pos;
      const double v1 = d[cnt - pos], v2 = d[cnt - pos - 1];
      v = v1 - (v1 - v2) * bias;
    }
  } else {
    // WeightedPercentileFun: float running sums in sorted order (sequential, as upstream accumulates them)
    const float* w = wsorted + off[l];
    float total = w[0];
    for (int i = 1; i < cnt; ++i) total = total + w[i];
    const float threshold = alpha_is_float ? static_cast<float>(total * static_cast<float>(alpha)) : static_cast<float>(total * alpha);
    // pos = upper_bound(cdf, threshold): the first i with cdf[i] > threshold
    float cdf = 0.0f, cdf_pos = 0.0f;
    int pos = cnt;
    for (int i = 0; i < cnt; ++i) {
      cdf = (i == 0) ? w[0] : cdf + w[i];
      if (cdf > threshold) { pos = i; cdf_pos = cdf; break; }
    }
    pos = min(pos, cnt - 1);
    if (pos == 0 || pos == cnt - 1) {
      v = d[pos];
    } else {
      const double v1 = d[pos - 1], v2 = d[pos];
      const float cdf_next = cdf_pos + w[pos + 1];
      if (cdf_next - cdf_pos >= 1.0f) v = static_cast<double>((threshold - cdf_pos) / (cdf_next - cdf_pos)) * (v2 - v1) + v1;
      else v = v2;
    }
  }
  leaf_value[l] = v;
  out[l] = v; out[L + l] = 1.0;
}
__global__ void k_set_leaf_values(double* __restrict__ leaf_value, const double* __restrict__ v, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) leaf_value[i] = v[i];
}

// ------------------------------------------------------------------------------------------------ K8 + K3 fused
// Applies the finished tree to the training scores by walking it with every row's bins (row-major records, streamed once)
// instead of going through the leaves' index ranges, and — for the row-wise objectives — computes the next iteration's
// gradients and their max |g|, max |h| in the same pass (OBJ = 1 binary, 2 L2, 0 scores only).
// The decisions are the partition kernel's: bin == nan_bin ? default_left : bin <= threshold.
// A thread keeps its row's record (WPR = 8 or 16 words) in a private, bank-staggered slice of shared memory, so a level costs two
// shared loads — the 8-byte node and the bin byte — and a handful of ALU instructions (picking the byte out of registers with
// selects was 25 % slower: the kernel is issue-bound); WPR = 0 reads the bins of wider records in place.
constexpr int kApplyThreads = 256;
constexpr int kApplyMaxFeatures = 1 << 14;      // node word: feature (14 bits) | nan bin (9) | threshold (8) | default_left (1)

// ScoreMix: score = (score * pre + leaf) * post.  Gradient boosting uses (1, 1); a random forest keeps the running mean of its
// trees in the scores (pre = iterations so far, post = 1 / (iterations + 1): RF::TrainOneIter's MultiplyScore / UpdateScore /
// MultiplyScore).  bias: added to every leaf value before it is used (RF: the tree carries the initial score; Tree::AddBias).
// shrink != 0: the leaf values are multiplied by lr with Tree::Shrinkage's round-to-zero.
// (struct ScoreMix: booster.h)
// ctrl == nullptr: an explicit tree of nl_explicit leaves whose arrays the host uploaded (DART re-adds earlier trees); otherwise
// the tree the learner has just finished (ctrl->done), once (ctrl->applied).
// CAT: the dataset has categorical features; a categorical node carries the marker 511 in its NaN-bin field and its decision is
// the bin's bit in the node's cat_bits (read through the read-only cache: such nodes are few).
template <int OBJ, int WPR, bool CAT = false>
__global__ void __launch_bounds__(kApplyThreads) k_bt_apply_tree(const uint8_t* __restrict__ rm, unsigned rm_stride,
                                                                 double* __restrict__ score, const float* __restrict__ label,
                                                                 const float* __restrict__ weight, float* __restrict__ grad,
                                                                 float* __restrict__ hess, TreeArraysDev T,
                                                                 const FeatureMetaDev* __restrict__ fmeta, BatchCtrl* __restrict__ ctrl,
                                                                 double lr, int n, double sigmoid, double lw_neg, double lw_pos,
                                                                 ScoreMix mix, int nl_explicit) {
  __shared__ double s_val[1024];
  __shared__ uint2 s_node[1024];
  int nl;
  if (ctrl != nullptr) {
    nl = ctrl->num_leaves;
    if (!ctrl->done || ctrl->abort || ctrl->applied || nl <= 1) return;
  } else {
    nl = nl_explicit;
  }
  for (int i = threadIdx.x; i < nl; i += blockDim.x) {
    double v = T.leaf_value[i];
    const double zt = (double)1e-35f;
    if (mix.shrink) { v *= lr; if (v >= -zt && v <= zt) v = 0.0; }
    if (mix.bias != 0.0) { v += mix.bias; if (v >= -zt && v <= zt) v = 0.0; }
    s_val[i] = v;
    if (i < nl - 1) {
      const int f = T.split_feature_inner[i];
      const int dt = T.decision_type[i];
      const FeatureMetaDev fm = fmeta[f];
      // the bin whose rows follow default_left: the NaN bin, or the zero bin under zero_as_missing
      const unsigned nan_bin = (CAT && (dt & 1)) ? 511u
                               : fm.wide_cols != 0 ? static_cast<unsigned>(fm.nan_local)      // a column of a wide feature (dataset.h)
                               : (fm.missing_type == kMissingNaN) ? static_cast<unsigned>(fm.num_bin - 1)
                               : (fm.missing_type == kMissingZero) ? static_cast<unsigned>(fm.default_bin) : 256u;
      uint2 nd;
      nd.x = static_cast<unsigned>(f) | (nan_bin << 14) | (static_cast<unsigned>(T.threshold_in_bin[i]) << 23) | ((dt & 2) ? (1u << 31) : 0u);
      nd.y = (static_cast<unsigned>(T.left_child[i]) & 0xffffu) | (static_cast<unsigned>(T.right_child[i]) << 16);
      s_node[i] = nd;
    }
  }
  __syncthreads();
  // stride WPR + 1 words: for a given word (or byte) offset the 32 lanes of a warp fall into 32 different banks
  __shared__ unsigned int s_rec[kApplyThreads * ((WPR == 0 ? 0 : WPR) + 1)];
  unsigned int* my_rec = s_rec + threadIdx.x * (WPR + 1);
  const unsigned char* my_bytes = reinterpret_cast<const unsigned char*>(my_rec);
  float mg = 0.0f, mh = 0.0f;
  for (long long row = static_cast<long long>(blockIdx.x) * kApplyThreads + threadIdx.x; row < n;
       row += static_cast<long long>(gridDim.x) * kApplyThreads) {
    const uint8_t* rec = rm + static_cast<size_t>(row) * rm_stride;
    if (WPR != 0) {      // the thread's record goes into its private, bank-staggered slice of shared memory: a byte-indexable register file
      #pragma unroll
      for (int c = 0; c < WPR / 4; ++c) {
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(rec) + c);
        my_rec[4 * c] = v.x; my_rec[4 * c + 1] = v.y; my_rec[4 * c + 2] = v.z; my_rec[4 * c + 3] = v.w;
      }
    }
    const double sc0 = score[row];
    int node = nl > 1 ? 0 : -1;      // a single-leaf tree (explicit trees only): every row is in leaf 0
    if (node >= 0) do {
      const uint2 nd = s_node[node];
      const int f = nd.x & 0x3fffu;
      const unsigned nan_bin = (nd.x >> 14) & 0x1ffu, thr = (nd.x >> 23) & 0xffu;
      const unsigned bin = WPR != 0 ? static_cast<unsigned>(my_bytes[f]) : static_cast<unsigned>(__ldg(rec + f));
      bool left = (bin == nan_bin) ? ((nd.x >> 31) != 0) : (bin <= thr);
      if (CAT && nan_bin == 511u) left = ((__ldg(T.cat_bits + node * 8 + (bin >> 5)) >> (bin & 31u)) & 1u) != 0u;
      node = left ? static_cast<int>(static_cast<short>(nd.y & 0xffffu)) : static_cast<int>(static_cast<short>(nd.y >> 16));
    } while (node >= 0);
    const double sc = (sc0 * mix.pre + s_val[~node]) * mix.post;      // (x * 1.0 + v) * 1.0 == x + v exactly
    score[row] = sc;
    if (OBJ != 0) {
      float g, h;
      const bool has_w = weight != nullptr;
      const float w = has_w ? weight[row] : 0.0f;
      if (OBJ == 1) GradBinaryRow(sc, label[row], has_w, w, sigmoid, lw_neg, lw_pos, &g, &h);
      else GradL2Row(sc, label[row], has_w, w, &g, &h);
      grad[row] = g; hess[row] = h;
      mg = fmaxf(mg, fabsf(g)); mh = fmaxf(mh, fabsf(h));
    }
  }
  if (OBJ != 0) {
    for (int o = 16; o > 0; o >>= 1) {
      mg = fmaxf(mg, __shfl_xor_sync(0xffffffffu, mg, o));
      mh = fmaxf(mh, __shfl_xor_sync(0xffffffffu, mh, o));
    }
    if ((threadIdx.x & 31) == 0) {
      atomicMax(&ctrl->max_abs_g_bits, __float_as_uint(mg));
      atomicMax(&ctrl->max_abs_h_bits, __float_as_uint(mh));
    }
  }
}
// marks the finished tree as applied (one thread, after k_bt_apply_tree): a second apply pass for the same tree does nothing
__global__ void k_bt_mark_applied(BatchCtrl* ctrl) { if (ctrl->done) ctrl->applied = 1; }

}  // namespace lgb
