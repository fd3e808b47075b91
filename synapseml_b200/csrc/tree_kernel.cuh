// Persistent tree kernel (single-GPU learner): ONE cooperative launch grows a whole leaf-wise tree.
//
// The multi-kernel path (hist -> find -> part_count -> part_scatter per split) spends most of a 31-leaf tree on fixed
// per-launch costs: launch + drain, re-reading descriptors through dependent global loads, clearing 128 KB of shared
// histograms per CTA (profiles/r01_*: ~20 us of a small leaf's histogram launch is overhead).  Here 148 CTAs stay
// resident, the shared histograms are cleared once, descriptors are re-read after grid-wide barriers, and the phases of a
// split are separated by cooperative-groups grid syncs instead of kernel boundaries:
//     root:   histogram | find | select
//     split:  partition count | partition scatter (+ clear work area, publish children) | histogram (smaller child) |
//             find (both children; sibling by subtraction) | reduce + select next leaf
// The phase bodies are the same device functions the stand-alone kernels use (HistRows, BlockFindBest, PartCountBody,
// PartScatterBody, SelectAndSplit).  State that changes during the kernel (index buffers, flags, work area, pool, leaf and
// control descriptors) is never read through the read-only (.nc) path; grid.sync() orders it.
// With more than one machine the exchange of histograms goes through NCCL between launches, so the data-parallel learner
// keeps the multi-kernel path.
#pragma once
#include <cooperative_groups.h>

#include "kernels.cuh"

namespace lgb {

namespace cg = cooperative_groups;

constexpr int kTreeThreads = 512;     // 2 split-search sub-blocks per CTA; <= 128 registers per thread

struct TreeKernelArgs {
  const uint8_t* rm; unsigned rm_stride;
  const uint8_t* cm; long long n_pad;
  const FeatureMetaDev* fmeta;
  const uint16_t* gh16; const int2* ghfix;
  int* idx0; int* idx1; int es;
  unsigned int* flags; int* part_counts;
  long long* work; long long* pool;
  LeafDev* leaves; TreeCtrl* ctrl;
  SplitDev* feat_best; unsigned char* splittable;
  TreeArraysDev T;
  LearnerParams P;
  int F, work_elems, rows_per_flush, n_local, n_global;
};

template <bool QUANT>
__device__ __forceinline__ void TreeFindPhase(const TreeKernelArgs& a, bool root) {
  // items = (child, feature); every 256-thread sub-block takes items round-robin
  TreeCtrl* ctrl = a.ctrl;
  const int F = a.F;
  const int nchild = root ? 1 : 2;
  const int sub = threadIdx.x >> 8, b = threadIdx.x & 255;
  const int nsub_total = gridDim.x * (kTreeThreads / kFindThreads);
  const double gscale = ctrl->grad_scale, hscale = ctrl->hess_scale;
  const size_t slot_stride = static_cast<size_t>(F) * kBins * 2;
  const int cur_leaf = ctrl->cur_leaf, parent_slot = ctrl->parent_slot, smaller = ctrl->smaller, larger = ctrl->larger;
  for (int item = blockIdx.x * (kTreeThreads / kFindThreads) + sub; item < nchild * F; item += nsub_total) {
    const int c = item / F, f = item - c * F;
    const FeatureMetaDev fm = a.fmeta[f];
    const int leaf = root ? 0 : (c == 0 ? smaller : larger);
    const LeafDev& Lf = a.leaves[leaf];
    const size_t off = (static_cast<size_t>(f) * kBins + b) * 2;
    longlong2 v = *reinterpret_cast<const longlong2*>(a.work + off);
    bool used = true;
    if (!root) {
      used = a.splittable[static_cast<size_t>(cur_leaf) * F + f] != 0;
      if (c == 1) {
        const longlong2 p = *reinterpret_cast<const longlong2*>(a.pool + static_cast<size_t>(parent_slot) * slot_stride + off);
        v.x = p.x - v.x; v.y = p.y - v.y;
      }
    }
    *reinterpret_cast<longlong2*>(a.pool + static_cast<size_t>(Lf.hist_slot) * slot_stride + off) = v;
    SplitDev* outp = a.feat_best + static_cast<size_t>(c) * F + f;
    unsigned char* spl = a.splittable + static_cast<size_t>(a.P.num_leaves + c) * F + f;   // scratch rows, copied by the reducer
    if (!used) {
      if (b == 0) { *outp = EmptySplit(f, fm.real_index); *spl = 0; }
    } else {
      long long tg, th;
      BlockFindBest<QUANT>(v.x, v.y, fm, f, root, Lf.int_g, Lf.int_h, Lf.sum_g, Lf.sum_h, Lf.global_count, a.P, gscale, hscale,
                           outp, spl, &tg, &th, sub);
      if (root && f == 0 && b == 0) {
        a.leaves[0].int_g = tg; a.leaves[0].int_h = th;
        a.leaves[0].sum_g = static_cast<double>(tg) * gscale; a.leaves[0].sum_h = static_cast<double>(th) * hscale;
      }
    }
    SubSync(sub);     // the sub-block's shared scratch is reused by its next item
  }
}

// CTA 0: per-feature results of the searched children -> each child's best split and splittable flags, then the next leaf.
__device__ __forceinline__ void TreeReduceAndSelect(const TreeKernelArgs& a, bool root) {
  __shared__ double s_gain[32];
  __shared__ int s_feat[32], s_idx[32];
  TreeCtrl* ctrl = a.ctrl;
  const int F = a.F;
  const int nchild = root ? 1 : 2;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  for (int c = 0; c < nchild; ++c) {
    const int leaf = root ? 0 : (c == 0 ? ctrl->smaller : ctrl->larger);
    const SplitDev* fb = a.feat_best + static_cast<size_t>(c) * F;
    double bg = KMINSCORE; int bf = -1, bi = -1;
    for (int j = threadIdx.x; j < F; j += blockDim.x) {
      const double g = fb[j].gain; const int rf = fb[j].real_feature;
      if (bi < 0 || SplitBetter(g, rf, bg, bf)) { bg = g; bf = rf; bi = j; }
    }
    for (int o = 16; o > 0; o >>= 1) {
      const double g = __shfl_xor_sync(0xffffffffu, bg, o);
      const int rf = __shfl_xor_sync(0xffffffffu, bf, o);
      const int i = __shfl_xor_sync(0xffffffffu, bi, o);
      if (i >= 0 && (bi < 0 || SplitBetter(g, rf, bg, bf))) { bg = g; bf = rf; bi = i; }
    }
    __syncthreads();
    if (lane == 0) { s_gain[warp] = bg; s_feat[warp] = bf; s_idx[warp] = bi; }
    __syncthreads();
    if (threadIdx.x == 0) {
      bg = s_gain[0]; bf = s_feat[0]; bi = s_idx[0];
      for (int w = 1; w < nw; ++w)
        if (s_idx[w] >= 0 && (bi < 0 || SplitBetter(s_gain[w], s_feat[w], bg, bf))) { bg = s_gain[w]; bf = s_feat[w]; bi = s_idx[w]; }
      a.leaves[leaf].best = fb[bi];
    }
    const unsigned char* src = a.splittable + static_cast<size_t>(a.P.num_leaves + c) * F;
    unsigned char* dst = a.splittable + static_cast<size_t>(leaf) * F;
    for (int j = threadIdx.x; j < F; j += blockDim.x) dst[j] = src[j];
  }
  __threadfence_block();
  __syncthreads();
  SelectAndSplit(a.leaves, ctrl, a.T, a.fmeta, a.P);
}

template <int R, int BPL, bool FLT, bool ROOT>
__device__ __forceinline__ void TreeHistPhase(const TreeKernelArgs& a, unsigned lane_addr, unsigned int* hsm) {
  constexpr int LPR = 32 / R;
  constexpr int FT = LPR * BPL;
  const int tiles = (a.F + FT - 1) / FT;
  const int row_ctas = max(1, static_cast<int>(gridDim.x) / tiles);
  const int tile = blockIdx.x / row_ctas, rcta = blockIdx.x - tile * row_ctas;
  if (tile >= tiles) return;
  const LeafDev& lf = a.leaves[ROOT ? 0 : a.ctrl->smaller];
  const int begin = lf.begin, count = lf.count, buf = lf.buf;
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    a.ctrl->hist_rows += count;
    a.ctrl->hist_launches += 1;
  }
  int rows_per_cta = (count + row_ctas - 1) / row_ctas;
  rows_per_cta = max((rows_per_cta + 31) & ~31, kHistMinRowsPerCta);
  const int lo = rcta * rows_per_cta;
  const int hi = min(count, lo + rows_per_cta);
  if (lo >= hi) return;
  const int lane = threadIdx.x & 31;
  const int q = lane % LPR;
  const bool kmask = (tile * FT + q * BPL) < a.F;
  const uint8_t* rmc = a.rm + static_cast<size_t>(tile) * FT + (kmask ? q : 0) * BPL;
  asm volatile("" : "+l"(rmc));
  const int* idx = ROOT ? nullptr : (buf == 0 ? a.idx0 : a.idx1);
  HistRows<R, BPL, FLT, ROOT>(rmc, a.rm_stride, a.gh16, a.ghfix, idx, begin, lo, hi, lane_addr, hsm, a.work, a.F, tile,
                              a.rows_per_flush, kmask);
}

template <int R, int BPL, bool FLT>
__global__ void __launch_bounds__(kTreeThreads, 1) k_tree(const TreeKernelArgs a) {
  extern __shared__ unsigned int hsm_raw[];
  cg::grid_group grid = cg::this_grid();
  const unsigned raw_addr = static_cast<unsigned>(__cvta_generic_to_shared(hsm_raw));
  const unsigned sbase = (raw_addr + kHistArrayBytes - 1) & ~(kHistArrayBytes - 1);
  unsigned int* hsm = hsm_raw + ((sbase - raw_addr) >> 2);
  const unsigned lane_addr = sbase + (threadIdx.x & 31) * 4;
  for (int i = threadIdx.x; i < 4 * kBins * 32; i += blockDim.x) hsm[i] = 0u;     // once: every flush leaves them cleared
  // ---- tree state (what k_tree_init + the memsets do in the multi-kernel path)
  {
    const int L = a.P.num_leaves;
    const int gtid = blockIdx.x * blockDim.x + threadIdx.x, gsz = gridDim.x * blockDim.x;
    TreeArraysDev T = a.T;
    for (int i = gtid; i < L; i += gsz) {
      LeafDev z;
      z.begin = 0; z.count = 0; z.buf = -1; z.global_count = 0; z.int_g = z.int_h = 0; z.sum_g = z.sum_h = 0.0;
      z.depth = 0; z.hist_slot = i;
      z.best = EmptySplit(-1, -1);
      if (i == 0) { z.count = a.n_local; z.global_count = a.n_global; }
      a.leaves[i] = z;
      T.leaf_parent[i] = -1; T.leaf_depth[i] = 0; T.leaf_value[i] = 0.0; T.leaf_weight[i] = 0.0; T.leaf_count[i] = 0;
      if (i < L - 1) {
        T.left_child[i] = 0; T.right_child[i] = 0; T.split_feature_inner[i] = 0; T.threshold_in_bin[i] = 0;
        T.internal_count[i] = 0; T.decision_type[i] = 0; T.split_gain[i] = 0.0f; T.internal_value[i] = 0.0; T.internal_weight[i] = 0.0;
      }
    }
    for (int i = gtid; i < a.work_elems; i += gsz) a.work[i] = 0;
    for (int i = gtid; i < (L + 2) * a.F; i += gsz) a.splittable[i] = 1;
    if (gtid == 0) {
      TreeCtrl* c = a.ctrl;
      c->num_leaves = 1; c->done = 0; c->pending = 1; c->find = 1; c->smaller = 0; c->larger = -1; c->cur_leaf = 0; c->new_leaf = 0;
      c->find_count[0] = c->find_count[1] = 0; c->find_children_done = 0;
    }
  }
  grid.sync();
  // phase timing (SM cycles seen by CTA 0; a phase ends at the grid barrier after it): the histogram share feeds the roofline
  // report, the rest is a development aid
  const bool timer = blockIdx.x == 0 && threadIdx.x == 0;
  long long t_tree = 0, t0 = 0, ph[6] = {0, 0, 0, 0, 0, 0};
  int dbg_i = 0;
#define PHASE_BEGIN() if (timer) t0 = clock64()
#define PHASE_END(k) if (timer) { const long long dt_ = clock64() - t0; ph[k] += dt_; if (dbg_i < 32) a.ctrl->dbg[dbg_i * 6 + k] = dt_; }
  if (a.P.seed == 4242u) {      // development aid: cost of an empty grid barrier
    long long tb = clock64();
    for (int i = 0; i < 10; ++i) grid.sync();
    if (timer) a.ctrl->dbg[191] = (clock64() - tb) / 10;
  }
  if (timer) t_tree = clock64();
  // ---- root
  PHASE_BEGIN();
  TreeHistPhase<R, BPL, FLT, true>(a, lane_addr, hsm);
  grid.sync();
  PHASE_END(1);
  PHASE_BEGIN();
  TreeFindPhase<!FLT>(a, true);
  grid.sync();
  PHASE_END(2);
  PHASE_BEGIN();
  if (blockIdx.x == 0) TreeReduceAndSelect(a, true);
  grid.sync();
  PHASE_END(3);
  // ---- splits
  const int num_splits = a.P.num_leaves - 1;
  for (int split = 0; split < num_splits; ++split) {
    if (*reinterpret_cast<volatile int*>(&a.ctrl->done)) break;     // same value on every CTA: written before the last grid.sync
    PHASE_BEGIN();
    PartCountBody(a.cm, a.n_pad, a.idx0, a.idx1, a.es, a.ctrl, a.flags, a.part_counts);
    grid.sync();
    PHASE_END(4);
    PHASE_BEGIN();
    PartScatterBody(a.idx0, a.idx1, a.idx0, a.idx1, a.es, a.gh16, a.ghfix, a.ctrl, a.leaves, a.flags, a.part_counts, a.T, a.work, a.work_elems, a.P);
    grid.sync();
    PHASE_END(5);
    if (split == num_splits - 1) break;
    const int find = *reinterpret_cast<volatile int*>(&a.ctrl->find);
    if (find) {
      PHASE_BEGIN();
      TreeHistPhase<R, BPL, FLT, false>(a, lane_addr, hsm);
      grid.sync();
      PHASE_END(1);
      PHASE_BEGIN();
      TreeFindPhase<!FLT>(a, false);
      grid.sync();
      PHASE_END(2);
    }
    PHASE_BEGIN();
    if (blockIdx.x == 0) {
      if (find) TreeReduceAndSelect(a, false);
      else SelectAndSplit(a.leaves, a.ctrl, a.T, a.fmeta, a.P);
    }
    grid.sync();
    PHASE_END(3);
    if (timer && dbg_i < 32) a.ctrl->dbg[dbg_i * 6 + 0] = a.leaves[a.ctrl->smaller].count;
    ++dbg_i;
  }
  if (timer) {
    for (int k = 0; k < 6; ++k) a.ctrl->phase_cycles[k] += ph[k];
    a.ctrl->hist_cycles += ph[1];
    a.ctrl->tree_cycles += clock64() - t_tree;
  }
#undef PHASE_BEGIN
#undef PHASE_END
}

}  // namespace lgb
