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

constexpr int kMaxBatch = 8;

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
  SplitDev best;
};

struct BatchItem {
  int node, left, right;
  int feature, threshold, default_left, nan_bin;
  int src_begin, src_count, src_buf;
  int smaller, larger, find;
  int parent_slot;
  int pad_[2];
};

struct BatchCtrl {
  int num_leaves;          // leaves of the committed tree
  int done;
  int num_nodes;           // nodes handed out so far
  int num_slots;           // histogram pool slots handed out so far
  int batch_n;
  int spec_used;           // expansions beyond the first of each batch so far (statistics)
  int rounds;
  int pad_;
  long long batch_rows;    // rows of all smaller children of the current batch (histogram CTAs are shared out by it)
  BatchItem batch[kMaxBatch];
  int find_count[kMaxBatch][2];
  unsigned int max_abs_g_bits, max_abs_h_bits;
  double grad_scale, hess_scale, inv_grad_scale, inv_hess_scale;
  long long hist_rows, hist_launches;
};

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
    z.best = EmptySplit(-1, -1);
    nodes[0] = z;
    avail[0] = 0;
    ctrl->num_leaves = 1; ctrl->done = 0; ctrl->num_nodes = 1; ctrl->num_slots = 1; ctrl->batch_n = 0; ctrl->spec_used = 0;
    ctrl->rounds = 0; ctrl->batch_rows = 0;
    for (int k = 0; k < kMaxBatch; ++k) { ctrl->find_count[k][0] = 0; ctrl->find_count[k][1] = 0; }
  }
}

// ------------------------------------------------------------------------------------------------ K4 wrapper
// grid = (row CTAs, feature tiles, batch slots).  The row CTAs of a round are shared out over the batch by row count.
template <int R, int BPL, bool FLT, bool ROOT>
__global__ void __launch_bounds__(kHistThreads, 1)
k_bt_hist(const uint8_t* __restrict__ rm, unsigned rm_stride, const uint16_t* __restrict__ gh16, const int2* __restrict__ ghfix,
          const int* __restrict__ idx0, const int* __restrict__ idx1, const NodeDev* __restrict__ nodes,
          BatchCtrl* __restrict__ ctrl, long long* __restrict__ work, int work_elems, int F, int rows_per_flush, int row_ctas) {
  extern __shared__ unsigned int hsm_raw[];
  constexpr int LPR = 32 / R;
  constexpr int FT = LPR * BPL;
  PdlLaunchDependents();
  const int lane = threadIdx.x & 31;
  const int q = lane % LPR;
  const int tile = blockIdx.y, k = blockIdx.z;
  const unsigned raw_addr = static_cast<unsigned>(__cvta_generic_to_shared(hsm_raw));
  const unsigned sbase = (raw_addr + kHistArrayBytes - 1) & ~(kHistArrayBytes - 1);
  unsigned int* hsm = hsm_raw + ((sbase - raw_addr) >> 2);
  const unsigned lane_addr = sbase + lane * 4;
  const bool kmask = (tile * FT + q * BPL) < F;
  const uint8_t* rmc = rm + static_cast<size_t>(tile) * FT + (kmask ? q : 0) * BPL;
  asm volatile("" : "+l"(rmc));
  PdlWait();
  if (ctrl->done) return;
  int node;
  long long total_rows;
  if (ROOT) { node = 0; total_rows = nodes[0].count; }
  else {
    if (k >= ctrl->batch_n || !ctrl->batch[k].find) return;
    node = ctrl->batch[k].smaller;
    total_rows = ctrl->batch_rows;
  }
  const NodeDev& nd = nodes[node];
  const int begin = nd.begin, count = nd.count, buf = nd.buf;
  if (blockIdx.x == 0 && tile == 0 && threadIdx.x == 0) {
    atomicAdd(reinterpret_cast<unsigned long long*>(&ctrl->hist_rows), static_cast<unsigned long long>(count));
    atomicAdd(reinterpret_cast<unsigned long long*>(&ctrl->hist_launches), 1ULL);
  }
  long long rpc = (total_rows + row_ctas - 1) / row_ctas;
  int rows_per_cta = static_cast<int>(rpc > 0x3fffffff ? 0x3fffffff : rpc);
  rows_per_cta = max((rows_per_cta + 31) & ~31, kHistMinRowsPerCta);
  const long long lo_ll = static_cast<long long>(blockIdx.x) * rows_per_cta;
  if (lo_ll >= count) return;
  const int lo = static_cast<int>(lo_ll);
  const int hi = min(count, lo + rows_per_cta);
  for (int i = threadIdx.x; i < 4 * kBins * 32; i += blockDim.x) hsm[i] = 0u;
  const int* idx = ROOT ? nullptr : (buf == 0 ? idx0 : idx1);
  __syncthreads();
  HistRows<R, BPL, FLT, ROOT>(rmc, rm_stride, gh16, ghfix, idx, begin, lo, hi, lane_addr, hsm,
                              work + static_cast<size_t>(ROOT ? 0 : k) * work_elems, F, tile, rows_per_flush, kmask);
}

// ------------------------------------------------------------------------------------------------ K5 + K6
// grid = (F, root ? 1 : 2, batch slots).  CTA (f, c, k): child c of batch item k (c = 0: the freshly built smaller child,
// c = 1: the larger child = parent - smaller).  The last CTA of each (k, c) reduces the per-feature results to the child's
// best split.
template <bool QUANT>
__global__ void __launch_bounds__(256) k_bt_find(const long long* __restrict__ work, int work_elems, long long* __restrict__ pool,
                                                 const FeatureMetaDev* __restrict__ fmeta, NodeDev* __restrict__ nodes,
                                                 BatchCtrl* __restrict__ ctrl, SplitDev* __restrict__ feat_best,
                                                 unsigned char* __restrict__ splittable, LearnerParams P, int root) {
  PdlLaunchDependents();
  PdlWait();
  if (ctrl->done) return;
  __shared__ int s_last;
  __shared__ double s_gain[8];
  __shared__ int s_feat[8], s_idx[8];
  const int F = P.num_features;
  const int f = blockIdx.x, c = blockIdx.y, k = blockIdx.z, b = threadIdx.x;
  int node, parent = -1, parent_slot = 0;
  if (root) node = 0;
  else {
    if (k >= ctrl->batch_n || !ctrl->batch[k].find) return;
    const BatchItem& it = ctrl->batch[k];
    node = c == 0 ? it.smaller : it.larger;
    parent = it.node;
    parent_slot = it.parent_slot;
  }
  const FeatureMetaDev fm = fmeta[f];
  const double gscale = ctrl->grad_scale, hscale = ctrl->hess_scale;
  const size_t slot_stride = static_cast<size_t>(F) * kBins * 2;
  const size_t off = (static_cast<size_t>(f) * kBins + b) * 2;
  const NodeDev& Nd = nodes[node];
  longlong2 v = *reinterpret_cast<const longlong2*>(work + static_cast<size_t>(root ? 0 : k) * work_elems + off);
  bool used = true;
  if (!root) {
    used = splittable[static_cast<size_t>(parent) * F + f] != 0;
    if (c == 1) {
      const longlong2 p = *reinterpret_cast<const longlong2*>(pool + static_cast<size_t>(parent_slot) * slot_stride + off);
      v.x = p.x - v.x; v.y = p.y - v.y;
    }
  }
  *reinterpret_cast<longlong2*>(pool + static_cast<size_t>(Nd.hist_slot) * slot_stride + off) = v;
  SplitDev* fb = feat_best + static_cast<size_t>(root ? 0 : (k * 2 + c)) * F;
  unsigned char* spl = splittable + static_cast<size_t>(node) * F + f;
  if (!used) {
    if (b == 0) { fb[f] = EmptySplit(f, fm.real_index); *spl = 0; }
  } else {
    long long tg, th;
    BlockFindBest<QUANT>(v.x, v.y, fm, f, root != 0, Nd.int_g, Nd.int_h, Nd.sum_g, Nd.sum_h, Nd.global_count, P, gscale, hscale,
                         fb + f, spl, &tg, &th, 0);
    if (root && f == 0 && b == 0) {
      nodes[0].int_g = tg; nodes[0].int_h = th;
      nodes[0].sum_g = static_cast<double>(tg) * gscale; nodes[0].sum_h = static_cast<double>(th) * hscale;
    }
  }
  __threadfence();
  __syncthreads();
  if (b == 0) s_last = (atomicAdd(&ctrl->find_count[root ? 0 : k][c], 1) == F - 1);
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  double bg = KMINSCORE; int bf = -1, bi = -1;
  for (int j = b; j < F; j += blockDim.x) {
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
  __syncthreads();
  if (b == 0) {
    bg = s_gain[0]; bf = s_feat[0]; bi = s_idx[0];
    for (int w = 1; w < 8; ++w)
      if (s_idx[w] >= 0 && (bi < 0 || SplitBetter(s_gain[w], s_feat[w], bg, bf))) { bg = s_gain[w]; bf = s_feat[w]; bi = s_idx[w]; }
    nodes[node].best = fb[bi];
    ctrl->find_count[root ? 0 : k][c] = 0;
  }
}

// ------------------------------------------------------------------------------------------------ K7 (batched)
__global__ void __launch_bounds__(kPartThreads) k_bt_part_count(const uint8_t* __restrict__ cm, long long n_pad,
                                                                const int* __restrict__ idx0, const int* __restrict__ idx1, int es,
                                                                const BatchCtrl* __restrict__ ctrl, unsigned int* __restrict__ flags,
                                                                long long flag_stride, int* __restrict__ block_counts) {
  PdlLaunchDependents();
  PdlWait();
  if (ctrl->done) return;
  const int k = blockIdx.y;
  if (k >= ctrl->batch_n) return;
  const BatchItem& it = ctrl->batch[k];
  const int n = it.src_count, begin = it.src_begin, buf = it.src_buf;
  const int thr = it.threshold, nan_bin = it.nan_bin, dl = it.default_left;
  const uint8_t* col = cm + static_cast<size_t>(it.feature) * n_pad;
  const int* idx = buf < 0 ? nullptr : (buf == 0 ? idx0 : idx1);
  flags += static_cast<size_t>(k) * flag_stride;
  block_counts += k * gridDim.x;
  int lo_w, hi_w;
  PartChunk(n, &lo_w, &hi_w);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  constexpr int NW = kPartThreads / 32;
  int cnt = 0;
  for (int w0 = lo_w + warp; w0 < hi_w; w0 += NW * 4) {
    int rows[4];
    #pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int w = w0 + u * NW;
      const int p = w * 32 + lane;
      rows[u] = (w < hi_w && p < n) ? (idx ? idx[static_cast<size_t>(begin + p) * es] : begin + p) : -1;
    }
    int bins[4];
    #pragma unroll
    for (int u = 0; u < 4; ++u) bins[u] = rows[u] >= 0 ? col[rows[u]] : -1;
    #pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int w = w0 + u * NW;
      const bool left = rows[u] >= 0 && ((bins[u] == nan_bin) ? (dl != 0) : (bins[u] <= thr));
      const unsigned int word = __ballot_sync(0xffffffffu, left);
      if (lane == 0 && w < hi_w) { flags[w] = word; cnt += __popc(word); }
    }
  }
  __shared__ int s_cnt[NW];
  if (lane == 0) s_cnt[warp] = cnt;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
    for (int i = 0; i < NW; ++i) t += s_cnt[i];
    block_counts[blockIdx.x] = t;
  }
}

__global__ void __launch_bounds__(kPartThreads) k_bt_part_scatter(const int* __restrict__ idx0r, const int* __restrict__ idx1r,
                                                                  int* __restrict__ idx0w, int* __restrict__ idx1w, int es,
                                                                  const uint16_t* __restrict__ gh16, const int2* __restrict__ ghfix,
                                                                  BatchCtrl* __restrict__ ctrl, NodeDev* __restrict__ nodes,
                                                                  const unsigned int* __restrict__ flags, long long flag_stride,
                                                                  const int* __restrict__ block_counts, long long* __restrict__ work,
                                                                  int work_elems, LearnerParams P) {
  PdlLaunchDependents();
  PdlWait();
  if (ctrl->done) return;
  const int k = blockIdx.y;
  if (k >= ctrl->batch_n) return;
  // the histogram of this item accumulates into work[k] with atomics: clear it here
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < work_elems; i += gridDim.x * blockDim.x)
    work[static_cast<size_t>(k) * work_elems + i] = 0;
  const BatchItem it = ctrl->batch[k];
  const int n = it.src_count, begin = it.src_begin, buf = it.src_buf;
  const int* idx = buf < 0 ? nullptr : (buf == 0 ? idx0r : idx1r);
  const int other = buf < 0 ? 0 : 1 - buf;
  int* dst = other == 0 ? idx0w : idx1w;
  flags += static_cast<size_t>(k) * flag_stride;
  block_counts += k * gridDim.x;
  constexpr int NW = kPartThreads / 32;
  __shared__ int s_red[2][NW];
  __shared__ int s_prefix[kPartWordsPerIter];
  __shared__ int s_carry;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int before = 0, total = 0;
  for (int i = threadIdx.x; i < gridDim.x; i += blockDim.x) {
    const int c = block_counts[i];
    total += c;
    if (i < blockIdx.x) before += c;
  }
  for (int o = 16; o > 0; o >>= 1) { before += __shfl_xor_sync(0xffffffffu, before, o); total += __shfl_xor_sync(0xffffffffu, total, o); }
  if (lane == 0) { s_red[0][warp] = before; s_red[1][warp] = total; }
  __syncthreads();
  before = 0; total = 0;
  for (int i = 0; i < NW; ++i) { before += s_red[0][i]; total += s_red[1][i]; }
  int lo_w, hi_w;
  PartChunk(n, &lo_w, &hi_w);
  if (threadIdx.x == 0) s_carry = before;
  __syncthreads();
  for (int base_w = lo_w; base_w < hi_w; base_w += kPartWordsPerIter) {
    const int nw = min(kPartWordsPerIter, hi_w - base_w);
    for (int t0 = 0; t0 < nw; t0 += kPartThreads) {
      const int t = t0 + threadIdx.x;
      int c = t < nw ? __popc(flags[base_w + t]) : 0;
      int incl = c;
      for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += v; }
      if (lane == 31) s_red[0][warp] = incl;
      __syncthreads();
      int wp = 0, tot = 0;
      for (int i = 0; i < NW; ++i) { if (i < warp) wp += s_red[0][i]; tot += s_red[0][i]; }
      if (t < nw) s_prefix[t] = s_carry + wp + incl - c;
      __syncthreads();
      if (threadIdx.x == 0) s_carry += tot;
      __syncthreads();
    }
    for (int w0 = warp; w0 < nw; w0 += NW * 4) {
      int4 ent[4];
      #pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int w = w0 + u * NW;
        const int p = (base_w + w) * 32 + lane;
        ent[u] = make_int4(-1, 0, 0, 0);
        if (w < nw && p < n) {
          if (idx) {
            if (es == 2) { const int2 e = reinterpret_cast<const int2*>(idx)[begin + p]; ent[u].x = e.x; ent[u].y = e.y; }
            else ent[u] = reinterpret_cast<const int4*>(idx)[begin + p];
          } else {           // the root: rows are 0..n-1, gradients come from the per-row arrays
            const int row = begin + p;
            ent[u].x = row;
            if (es == 2) {
              const unsigned int g = gh16[row];
              ent[u].y = static_cast<int>((static_cast<unsigned>(static_cast<int>(static_cast<signed char>(g >> 8))) << 16) + (g & 0xffu));
            } else { const int2 g = ghfix[row]; ent[u].y = g.x; ent[u].z = g.y; }
          }
        }
      }
      #pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int w = w0 + u * NW;
        if (w >= nw) continue;
        const unsigned int word = flags[base_w + w];
        const int p = (base_w + w) * 32 + lane;
        if (ent[u].x >= 0) {
          const int lefts_before = s_prefix[w] + __popc(word & ((1u << lane) - 1u));
          const int pos = ((word >> lane) & 1u) ? begin + lefts_before : begin + total + (p - lefts_before);
          if (es == 2) reinterpret_cast<int2*>(dst)[pos] = make_int2(ent[u].x, ent[u].y);
          else reinterpret_cast<int4*>(dst)[pos] = ent[u];
        }
      }
    }
    __syncthreads();
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    // publish the two children (what DataPartition::Split + the leaf-splits initialisation do)
    const NodeDev par = nodes[it.node];
    const SplitDev bs = par.best;
    const int nl_true = total, nr_true = n - total;
    int lc = bs.left_count, rc = bs.right_count;
    if (!P.dp) { lc = nl_true; rc = nr_true; }       // serial learner: true partition counts replace the estimates
    NodeDev L, Rr;
    const int depth = par.depth + 1;
    L.begin = begin; L.count = nl_true; L.buf = other; L.global_count = lc;
    L.int_g = bs.left_int_g; L.int_h = bs.left_int_h; L.sum_g = bs.left_sum_g; L.sum_h = bs.left_sum_h; L.depth = depth;
    Rr.begin = begin + nl_true; Rr.count = nr_true; Rr.buf = other; Rr.global_count = rc;
    Rr.int_g = bs.right_int_g; Rr.int_h = bs.right_int_h; Rr.sum_g = bs.right_sum_g; Rr.sum_h = bs.right_sum_h; Rr.depth = depth;
    L.state = Rr.state = 0; L.left = L.right = Rr.left = Rr.right = -1; L.pad_ = Rr.pad_ = 0;
    L.best = EmptySplit(-1, -1); Rr.best = EmptySplit(-1, -1);
    const bool left_smaller = lc < rc;
    const int fresh_slot = atomicAdd(&ctrl->num_slots, 1);
    // the larger child takes over the parent's pool slot (parent - smaller is formed in place), the smaller gets a fresh one
    L.hist_slot = left_smaller ? fresh_slot : par.hist_slot;
    Rr.hist_slot = left_smaller ? par.hist_slot : fresh_slot;
    bool find = true;                                  // BeforeFindBestSplit
    if (P.max_depth > 0 && depth >= P.max_depth) find = false;
    if (rc < P.min_data_in_leaf * 2 && lc < P.min_data_in_leaf * 2) find = false;
    nodes[it.left] = L; nodes[it.right] = Rr;
    nodes[it.node].state = 1; nodes[it.node].left = it.left; nodes[it.node].right = it.right;
    ctrl->batch[k].smaller = left_smaller ? it.left : it.right;
    ctrl->batch[k].larger = left_smaller ? it.right : it.left;
    ctrl->batch[k].find = find ? 1 : 0;
    if (find) atomicAdd(reinterpret_cast<unsigned long long*>(&ctrl->batch_rows), static_cast<unsigned long long>(left_smaller ? nl_true : nr_true));
  }
}

// ------------------------------------------------------------------------------------------------ replay + next batch
__global__ void __launch_bounds__(256) k_bt_select(NodeDev* __restrict__ nodes, int* __restrict__ avail, BatchCtrl* __restrict__ ctrl,
                                                   TreeArraysDev T, const FeatureMetaDev* __restrict__ fmeta, LearnerParams P,
                                                   int max_nodes, int spec_budget, int max_batch) {
  PdlLaunchDependents();
  PdlWait();
  if (ctrl->done) return;
  __shared__ double s_gain[1024];
  __shared__ int s_feat[1024], s_node[1024], s_rank[1024];
  __shared__ unsigned char s_exp[1024];
  __shared__ double r_gain[8];
  __shared__ int r_feat[8], r_idx[8];
  __shared__ int s_best, s_state;      // s_state: 0 keep replaying, 1 finished tree, 2 expansion needed
  __shared__ int s_count;
  const int L = P.num_leaves;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  int nl = ctrl->num_leaves;
  for (int i = tid; i < nl; i += blockDim.x) {
    const int nd = avail[i];
    s_node[i] = nd; s_gain[i] = nodes[nd].best.gain; s_feat[i] = nodes[nd].best.real_feature; s_exp[i] = nodes[nd].state == 1;
  }
  __syncthreads();
  // ---- replay the sequential order as far as the already expanded leaves allow
  while (true) {
    double bg = KMINSCORE; int bf = -1, bi = -1;
    for (int i = tid; i < nl; i += blockDim.x) {
      const double g = s_gain[i]; const int rf = s_feat[i];
      if (bi < 0 || SplitBetter(g, rf, bg, bf)) { bg = g; bf = rf; bi = i; }        // ascending i: the earlier leaf keeps ties
    }
    auto take = [&](double g, int rf, int i) {
      if (i < 0) return;
      if (bi < 0 || SplitBetter(g, rf, bg, bf) || (!SplitBetter(bg, bf, g, rf) && i < bi)) { bg = g; bf = rf; bi = i; }
    };
    for (int o = 16; o > 0; o >>= 1) {
      const double g = __shfl_xor_sync(0xffffffffu, bg, o);
      const int rf = __shfl_xor_sync(0xffffffffu, bf, o);
      const int i = __shfl_xor_sync(0xffffffffu, bi, o);
      take(g, rf, i);
    }
    if (lane == 0) { r_gain[warp] = bg; r_feat[warp] = bf; r_idx[warp] = bi; }
    __syncthreads();
    if (tid == 0) {
      bg = r_gain[0]; bf = r_feat[0]; bi = r_idx[0];
      for (int w = 1; w < 8; ++w) take(r_gain[w], r_feat[w], r_idx[w]);
      s_best = bi;
      if (nl >= L || !(bg > 0.0)) s_state = 1;
      else if (!s_exp[bi]) s_state = 2;
      else {
        // ---- commit: Tree::Split with the leaf ids of the sequential learner (left keeps the id, right = num_leaves)
        s_state = 0;
        const int best_leaf = bi, nd = s_node[bi];
        const NodeDev par = nodes[nd];
        const SplitDev bs = par.best;
        const int lnode = par.left, rnode = par.right;
        const int lc = nodes[lnode].global_count, rc = nodes[rnode].global_count;
        const int new_node = nl - 1, new_leaf = nl;
        const FeatureMetaDev fm = fmeta[bs.feature];
        const int parent = T.leaf_parent[best_leaf];
        if (parent >= 0) {
          if (T.left_child[parent] == ~best_leaf) T.left_child[parent] = new_node; else T.right_child[parent] = new_node;
        }
        T.split_feature_inner[new_node] = bs.feature;
        T.split_gain[new_node] = static_cast<float>(bs.gain);
        T.left_child[new_node] = ~best_leaf;
        T.right_child[new_node] = ~new_leaf;
        T.leaf_parent[best_leaf] = new_node;
        T.leaf_parent[new_leaf] = new_node;
        T.internal_weight[new_node] = bs.left_sum_h + bs.right_sum_h;
        T.internal_value[new_node] = T.leaf_value[best_leaf];
        T.internal_count[new_node] = lc + rc;
        T.leaf_value[best_leaf] = (bs.left_output != bs.left_output) ? 0.0 : bs.left_output;
        T.leaf_weight[best_leaf] = bs.left_sum_h;
        T.leaf_count[best_leaf] = lc;
        T.leaf_value[new_leaf] = (bs.right_output != bs.right_output) ? 0.0 : bs.right_output;
        T.leaf_weight[new_leaf] = bs.right_sum_h;
        T.leaf_count[new_leaf] = rc;
        T.leaf_depth[new_leaf] = T.leaf_depth[best_leaf] + 1;
        T.leaf_depth[best_leaf]++;
        signed char dt = 0;
        if (bs.default_left) dt |= 2;
        dt |= static_cast<signed char>((fm.missing_type & 3) << 2);
        T.decision_type[new_node] = dt;
        T.threshold_in_bin[new_node] = bs.threshold;
        nodes[nd].state = 2;
        avail[best_leaf] = lnode; avail[new_leaf] = rnode;
        s_node[best_leaf] = lnode; s_gain[best_leaf] = nodes[lnode].best.gain; s_feat[best_leaf] = nodes[lnode].best.real_feature;
        s_exp[best_leaf] = nodes[lnode].state == 1;
        s_node[new_leaf] = rnode; s_gain[new_leaf] = nodes[rnode].best.gain; s_feat[new_leaf] = nodes[rnode].best.real_feature;
        s_exp[new_leaf] = nodes[rnode].state == 1;
      }
    }
    __syncthreads();
    if (s_state != 0) break;
    ++nl;
    __syncthreads();
  }
  if (s_state == 1) {
    if (tid == 0) { ctrl->done = 1; ctrl->num_leaves = nl; ctrl->batch_n = 0; }
    return;
  }
  // ---- next batch: the best not-yet-expanded leaves whose rank is below the remaining leaf budget
  const int budget = L - nl;
  for (int i = tid; i < nl; i += blockDim.x) {
    const double g = s_gain[i]; const int rf = s_feat[i];
    int rank = 0;
    for (int j = 0; j < nl; ++j) {
      if (j == i) continue;
      const double gj = s_gain[j]; const int fj = s_feat[j];
      rank += SplitBetter(gj, fj, g, rf) || (!SplitBetter(g, rf, gj, fj) && j < i);
    }
    s_rank[i] = rank;
  }
  if (tid == 0) s_count = 0;
  __syncthreads();
  // expansions that are not committed yet (the batch's first leaf is certain to be committed by the next replay, the others
  // are speculative): bounding them bounds the nodes / pool slots a tree can consume to (L - 1) + spec_budget expansions
  const int outstanding = (ctrl->num_nodes - 1) / 2 - (nl - 1);
  const int spec_left = max(0, spec_budget - outstanding);
  const int node_room = (max_nodes - ctrl->num_nodes) / 2;
  int kmax = min(min(max_batch, 1 + spec_left), node_room);
  if (kmax < 1) kmax = 1;
  for (int i = tid; i < nl; i += blockDim.x) {
    const bool eligible = !s_exp[i] && s_gain[i] > 0.0 && s_rank[i] < budget;
    if (!eligible) continue;
    int pos = 0;
    for (int j = 0; j < nl; ++j)
      if (j != i && !s_exp[j] && s_gain[j] > 0.0 && s_rank[j] < budget && s_rank[j] < s_rank[i]) ++pos;
    if (pos < kmax) {
      atomicAdd(&s_count, 1);
      const int nd = s_node[i];
      const NodeDev par = nodes[nd];
      const SplitDev bs = par.best;
      const FeatureMetaDev fm = fmeta[bs.feature];
      BatchItem it;
      it.node = nd; it.left = ctrl->num_nodes + 2 * pos; it.right = it.left + 1;
      it.feature = bs.feature; it.threshold = bs.threshold; it.default_left = bs.default_left;
      it.nan_bin = (fm.missing_type == kMissingNaN) ? fm.num_bin - 1 : -1;
      it.src_begin = par.begin; it.src_count = par.count; it.src_buf = par.buf;
      it.smaller = it.larger = -1; it.find = 0; it.parent_slot = par.hist_slot; it.pad_[0] = it.pad_[1] = 0;
      ctrl->batch[pos] = it;
    }
  }
  __syncthreads();
  if (tid == 0) {
    const int bn = s_count;
    ctrl->batch_n = bn;
    ctrl->num_nodes += 2 * bn;
    ctrl->spec_used += bn - 1;      // statistics only
    ctrl->num_leaves = nl;
    ctrl->batch_rows = 0;
    ctrl->rounds += 1;
  }
}

// ------------------------------------------------------------------------------------------------ K8 (batched learner)
__global__ void __launch_bounds__(256) k_bt_sort_leaves(const NodeDev* __restrict__ nodes, const int* __restrict__ avail,
                                                        const BatchCtrl* __restrict__ ctrl, int* __restrict__ sorted, int L) {
  const int nl = ctrl->num_leaves;
  for (int i = threadIdx.x; i < nl; i += blockDim.x) {
    const NodeDev& ni = nodes[avail[i]];
    const int bi = ni.begin, ci = ni.count;
    int rank = 0;
    for (int j = 0; j < nl; ++j) {
      const NodeDev& nj = nodes[avail[j]];
      const int bj = nj.begin, cj = nj.count;
      rank += (bj < bi) || (bj == bi && (cj < ci || (cj == ci && j < i)));
    }
    sorted[rank] = i;
    sorted[L + rank] = bi;
    sorted[2 * L + rank] = ni.buf;
  }
}

__global__ void __launch_bounds__(256) k_bt_score_update(double* __restrict__ score, const int* __restrict__ idx0,
                                                         const int* __restrict__ idx1, int es, const BatchCtrl* __restrict__ ctrl,
                                                         const int* __restrict__ sorted, int L, const double* __restrict__ leaf_value,
                                                         double lr, int n) {
  __shared__ int s_begin[1024], s_buf[1024];
  __shared__ double s_val[1024];
  const int nl = ctrl->num_leaves;
  if (nl <= 1) return;
  for (int i = threadIdx.x; i < nl; i += blockDim.x) {
    const int leaf = sorted[i];
    s_begin[i] = sorted[L + i]; s_buf[i] = sorted[2 * L + i];
    double v = leaf_value[leaf] * lr;
    const double zt = (double)1e-35f;
    if (v >= -zt && v <= zt) v = 0.0;
    s_val[i] = v;
  }
  __syncthreads();
  for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < n; p += gridDim.x * blockDim.x) {
    int lo = 0, hi = nl - 1;
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (s_begin[mid] <= p) lo = mid; else hi = mid - 1;
    }
    const int row = (s_buf[lo] == 0 ? idx0 : idx1)[static_cast<size_t>(p) * es];
    score[row] += s_val[lo];
  }
}

}  // namespace lgb
