// Persistent tree kernel (single-GPU learner): ONE cooperative launch grows a whole tree.
//
// A late tree of a long boosting run is deep and narrow: ~18 dependent rounds whose leaves are small, so the multi-kernel
// path (partition -> histogram -> split search -> select, 4 launches per round) spends most of its time on what surrounds
// the work — grid launch and drain, re-reading the control block through dependent loads, clearing 128 KB of shared
// histograms per CTA.  Here one CTA per SM stays resident for the whole tree, the shared histograms are cleared once, and the
// phases of a round are separated by cooperative-groups grid syncs (~1.3 us each, experiments/grid_sync_probe.py) instead of
// kernel boundaries:
//     root :  histogram | split search | select
//     round:  partition (+ clear work areas) | histogram of the smaller children |
//             split search of both children (sibling by subtraction; last unit of a child reduces over features) | select
// The phase bodies are the device functions the stand-alone kernels use (PartitionTiles, HistBody, FindUnit, SelectBody), so
// both paths produce the same trees.  Everything that changes during the kernel (index buffers, work areas, pool, nodes,
// control block) is read with plain loads; grid.sync() orders it.  The row records, bin columns and gradient arrays are
// constant during a tree and may use the read-only path.
// With more than one machine the histograms are exchanged through NCCL between launches, so the data-parallel learner keeps
// the multi-kernel path.
#pragma once
#include <cooperative_groups.h>

#include "batch_tree.cuh"

namespace lgb {

namespace cg = cooperative_groups;

struct TreeKernelArgs {
  const uint8_t* rm; unsigned rm_stride;
  const uint8_t* cm; long long n_pad;
  const FeatureMetaDev* fmeta;
  const uint16_t* gh16; const int2* ghfix;
  int* idx0; int* idx1;
  long long* work; long long* pool;
  NodeDev* nodes; int* avail; BatchCtrl* ctrl;
  SplitDev* feat_best; unsigned char* splittable;
  TreeArraysDev T;
  LearnerParams P;
  int F, work_elems, rows_per_flush, tiles, row_ctas;
  int max_nodes, spec_budget, max_batch;
};

constexpr int kTreeSubs = kHistThreads / kFindThreads;      // 256-thread sub-blocks per CTA

template <bool QUANT>
__device__ __forceinline__ void TreeFindPhase(const TreeKernelArgs& a, bool root, int* s_units /*[kMaxBatch + 1]*/) {
  // units = (batch slot, child, feature); every sub-block of the grid takes units round-robin
  BatchCtrl* ctrl = a.ctrl;
  const int F = a.F;
  const int sub = threadIdx.x >> 8;
  if (threadIdx.x == 0) {
    int u = 0;
    const int bn = root ? 1 : ctrl->batch_n;
    for (int k = 0; k < kMaxBatch; ++k) {
      s_units[k] = u;
      if (k < bn) u += (root ? 1 : 2) * F;
    }
    s_units[kMaxBatch] = u;
  }
  __syncthreads();
  const int total = s_units[kMaxBatch];
  const int nsub = gridDim.x * kTreeSubs;
  for (int unit = blockIdx.x * kTreeSubs + sub; unit < total; unit += nsub) {
    int k = 0;
    while (k + 1 < kMaxBatch && unit >= s_units[k + 1]) ++k;
    const int r = unit - s_units[k];
    const int c = r / F, f = r - c * F;
    FindUnit<QUANT>(a.work, a.work_elems, a.pool, a.fmeta, a.nodes, ctrl, a.feat_best, a.splittable, a.P, root ? 1 : 0, f, c, k, sub);
  }
  __syncthreads();
}

template <int R, int BPL, bool FLT>
__global__ void __launch_bounds__(kHistThreads, 1) k_bt_tree(const TreeKernelArgs a) {
  extern __shared__ unsigned int hsm_raw[];
  __shared__ HistShared s_hist;
  __shared__ int s_list[kMaxBatch + 1];
  constexpr int ES = FLT ? 4 : 2;
  cg::grid_group grid = cg::this_grid();
  BatchCtrl* ctrl = a.ctrl;
  const unsigned raw_addr = static_cast<unsigned>(__cvta_generic_to_shared(hsm_raw));
  const unsigned sbase = (raw_addr + kHistArrayBytes - 1) & ~(kHistArrayBytes - 1);
  unsigned int* hsm = hsm_raw + ((sbase - raw_addr) >> 2);
  const unsigned lane_addr = sbase + (threadIdx.x & 31) * 4;
  double* sel_smem = reinterpret_cast<double*>(hsm + 4 * kBins * 32);      // after the four histogram arrays
  for (int i = threadIdx.x; i < 4 * kBins * 32; i += blockDim.x) hsm[i] = 0u;
  __syncthreads();
  const int tile = blockIdx.x / a.row_ctas, cta_x = blockIdx.x - tile * a.row_ctas;
  const bool hist_cta = tile < a.tiles;
  const int sub = threadIdx.x >> 8;
  // ---- root
  if (hist_cta)
    HistBody<R, BPL, FLT, true>(a.rm, a.rm_stride, a.gh16, a.ghfix, a.idx0, a.idx1, a.nodes, ctrl, a.work, a.work_elems, a.F,
                                a.rows_per_flush, a.row_ctas, cta_x, tile, hsm, lane_addr, &s_hist, a.P);
  grid.sync();
  TreeFindPhase<!FLT>(a, true, s_list);
  grid.sync();
  if (blockIdx.x == 0 && threadIdx.x < kFindThreads)
    SelectBody(a.nodes, a.avail, ctrl, a.T, a.fmeta, a.P, a.max_nodes, a.spec_budget, a.max_batch, sel_smem, 0, a.feat_best, 1);
  grid.sync();
  // ---- rounds
  while (!ctrl->done) {
    const int bn = ctrl->batch_n;
    for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < static_cast<size_t>(bn) * a.work_elems;
         i += static_cast<size_t>(gridDim.x) * blockDim.x)
      a.work[i] = 0;
    if (threadIdx.x == 0) PartTileList(ctrl, bn, s_list);
    __syncthreads();
    PartitionTiles<ES>(a.cm, a.n_pad, a.idx0, a.idx1, a.idx0, a.idx1, a.gh16, a.ghfix, ctrl, bn, s_list, blockIdx.x * kTreeSubs + sub,
                       gridDim.x * kTreeSubs, sub);
    grid.sync();
    if (hist_cta)
      HistBody<R, BPL, FLT, false>(a.rm, a.rm_stride, a.gh16, a.ghfix, a.idx0, a.idx1, a.nodes, ctrl, a.work, a.work_elems, a.F,
                                   a.rows_per_flush, a.row_ctas, cta_x, tile, hsm, lane_addr, &s_hist, a.P);
    grid.sync();
    TreeFindPhase<!FLT>(a, false, s_list);
    grid.sync();
    if (blockIdx.x == 0 && threadIdx.x < kFindThreads)
      SelectBody(a.nodes, a.avail, ctrl, a.T, a.fmeta, a.P, a.max_nodes, a.spec_budget, a.max_batch, sel_smem, 0, a.feat_best, 0);
    grid.sync();
  }
}

}  // namespace lgb
