// Histogram construction kernel (K4), in a header of its own so that more than one translation unit can instantiate it
// (booster.cu for training, noalloc.cu for the allocation-free entry point).  Everything here is a template or inline.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "booster.h"

namespace lgb {

constexpr int kBins = 256;
constexpr int kHistThreads = 1024;
constexpr int kHistMinRowsPerCta = 256;      // measured on C2, 500 iterations: 8192 -> 1.230, 2048 -> 1.191, 512 -> 1.179, 256 -> 1.176 ms per iteration

// Programmatic dependent launch (sm_90+): a kernel launched with the programmatic-stream-serialization attribute may start
// while its predecessor in the stream is still draining.  PdlLaunchDependents() lets the successor's CTAs become resident;
// PdlWait() blocks until the predecessor grid has completed and its writes are visible.  Everything before PdlWait() must
// not touch memory the predecessor writes.
__device__ __forceinline__ void PdlLaunchDependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void PdlWait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// ------------------------------------------------------------------------------------------------ K4 histogram
// Shared-memory layout [array][bin][32 lanes]: in any one warp-wide atomic every lane hits its own bank, whatever the
// bins are, so throughput does not depend on the bin distribution (measured on B200: 2.0-4.2 T updates/s against
// 0.7-3.5 T/s for a [feature][bin] layout; profiles/r01_smem_atomics_microbench.log).  A lane covers BPL consecutive
// feature bytes of one row; 32/R lanes cover one row record slice and R rows are handled per warp step, so each
// feature has R private sub-histograms that are summed when the CTA flushes to the global int64 histogram.
//   quantized: one RED.SHARED.ADD.U32 per (row, feature) on a packed (grad << 16) + hess word; a CTA flushes before any
//              16-bit lane can overflow.
//   float    : two RED.SHARED.ADD.U32 per (row, feature) on 32-bit fixed-point sums.
// The four 32 KB arrays start at a 32 KB-aligned shared address, so the address of an update is
// ((bins >> s) & 0x7f80) | lane_address — two ALU instructions per atomic.
// Algorithmic bytes per row visit: F * 1 (bins) + 2 (quantized g,h) or 8 (float) + 4 (row index; 0 at the root).
constexpr unsigned kHistArrayBytes = kBins * 32 * 4;       // 32 KB
constexpr size_t kHistSmemBytes = 4 * kHistArrayBytes + kHistArrayBytes;   // + alignment slack

// All of a lane's BPL updates for one row in one asm block: no per-update branch, array offsets as immediates.
template <int BPL, bool FLT>
__device__ __forceinline__ void HistUpdate(unsigned bins, unsigned lane_addr, unsigned pk, unsigned gx, unsigned hx) {
  // byte k of `bins` -> byte offset bin * 128 inside array k
  const unsigned a0 = ((bins << 7) & 0x7f80u) | lane_addr;
  const unsigned a1 = ((bins >> 1) & 0x7f80u) | lane_addr;
  if (FLT) {
    asm volatile(
        "red.shared.add.u32 [%0], %2;\n\t"
        "red.shared.add.u32 [%0+32768], %3;\n\t"
        "red.shared.add.u32 [%1+65536], %2;\n\t"
        "red.shared.add.u32 [%1+98304], %3;" ::"r"(a0), "r"(a1), "r"(gx), "r"(hx) : "memory");
  } else {
    const unsigned a2 = ((bins >> 9) & 0x7f80u) | lane_addr;
    const unsigned a3 = ((bins >> 17) & 0x7f80u) | lane_addr;
    asm volatile(
        "red.shared.add.u32 [%0], %4;\n\t"
        "red.shared.add.u32 [%1+32768], %4;\n\t"
        "red.shared.add.u32 [%2+65536], %4;\n\t"
        "red.shared.add.u32 [%3+98304], %4;" ::"r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(pk) : "memory");
  }
}

// Flush of a CTA's shared histograms (all threads; the caller synchronises before and after): sums the R sub-histograms of each
// (feature, bin), adds the sum into the global int64 histogram `work` and leaves the shared arrays zero.
// PACK (quantized mode only): `work` holds ONE int64 per (feature, bin), (sum g << 32) + sum h — valid while the global sum of
// h stays below 2^32 — so a flush is one global atomic per bin instead of two.
// Measured (experiments/flush_probe.cu, profiles/r02_flush_probe.log): with 148 CTAs adding into the same 7168 words the flush
// costs 5.2 us, 1.8 us of it the scan itself; reducing the CTAs of a cluster through distributed shared memory first is slower
// (8.3 / 16.5 / 24.5 us for clusters of 2 / 4 / 8: DSMEM moves ~20 B per cycle per SM), as are per-CTA slabs + a sum kernel.
template <int R, int BPL, bool FLT, bool PACK>
__device__ __forceinline__ void HistFlush(unsigned int* hsm, long long* work, int F, int tile) {
  constexpr int LPR = 32 / R;
  constexpr int FT = LPR * BPL;
  for (int i = threadIdx.x; i < BPL * kBins * LPR; i += blockDim.x) {
    const int qq = i % LPR;
    const int bin = (i / LPR) % kBins;
    const int k = i / (LPR * kBins);
    const int f = tile * FT + qq * BPL + k;
    long long g = 0, h = 0;
    // a warp covers R consecutive bins x LPR lanes; all bins share the banks of a given (sub-histogram, lane), so each thread
    // starts at a different sub-histogram ((bin + rr) mod R): 32 distinct banks per access instead of an R-way conflict
    const int skew = (i / LPR) & (R - 1);
    #pragma unroll
    for (int rr = 0; rr < R; ++rr) {
      const int a = bin * 32 + ((rr + skew) & (R - 1)) * LPR + qq;
      if (FLT) {
        g += static_cast<int>(hsm[(2 * k) * (kBins * 32) + a]);
        h += static_cast<int>(hsm[(2 * k + 1) * (kBins * 32) + a]);
        hsm[(2 * k) * (kBins * 32) + a] = 0u;
        hsm[(2 * k + 1) * (kBins * 32) + a] = 0u;
      } else {
        const int w = static_cast<int>(hsm[k * (kBins * 32) + a]);
        h += w & 0xffff;
        g += w >> 16;
        hsm[k * (kBins * 32) + a] = 0u;
      }
    }
    if (f < F && (g != 0 || h != 0)) {
      if (PACK) {
        atomicAdd(reinterpret_cast<unsigned long long*>(work + static_cast<size_t>(f) * kBins + bin),
                  static_cast<unsigned long long>(g * 4294967296LL + h));
      } else {
        unsigned long long* dst = reinterpret_cast<unsigned long long*>(work + (static_cast<size_t>(f) * kBins + bin) * 2);
        if (g != 0) atomicAdd(dst, static_cast<unsigned long long>(g));
        if (h != 0) atomicAdd(dst + 1, static_cast<unsigned long long>(h));
      }
    }
  }
}

// Accumulates rows [lo, hi) of a leaf (positions into its index range; ROOT: positions are row ids) into the CTA's shared
// histograms and flushes them into the global int64 histogram `work` (every rows_per_flush rows and at the end).  The shared
// arrays must be zero on entry and are zero again on return.  Shared by k_hist and the persistent tree kernel.
template <int R, int BPL, bool FLT, bool ROOT, bool PACK = false>
__device__ __forceinline__ void HistRows(const uint8_t* rmc, unsigned rm_stride, const uint16_t* __restrict__ gh16,
                                         const int2* __restrict__ ghfix, const int* idx, int begin, int lo, int hi,
                                         unsigned lane_addr, unsigned int* hsm, long long* work, int F, int tile,
                                         int rows_per_flush, bool kmask, unsigned long long* t_flush = nullptr) {
  constexpr int LPR = 32 / R;
  constexpr int FT = LPR * BPL;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  const int r = lane / LPR;
  constexpr int STEPS = 32 / R;
  constexpr int U = 8;
  for (int sub = lo; sub < hi; sub += rows_per_flush) {
    const int sub_hi = min(hi, sub + rows_per_flush);
    // ---- this warp's full 32-row groups, software-pipelined: while one unit's updates issue, the next unit's bins and the
    //      next group's row ids / gradients are already in flight (the kernel is latency-bound otherwise)
    const int first = sub + warp * 32;
    const int gstride = nwarps * 32;
    const int ng = (first + 32 <= sub_hi) ? (sub_hi - 32 - first) / gstride + 1 : 0;
    constexpr int NU = STEPS / U;
    // one index entry per row of a non-root leaf: {row, packed g|h} (quantized) or {row, g, h, -} (float); the root reads the
    // per-row gradient arrays directly (rows are consecutive)
    auto load_entry = [&](int p, int& row, unsigned& pk, unsigned& gx, unsigned& hx) {
      if (ROOT) {
        row = begin + p;
        if (FLT) { const int2 t = ghfix[row]; gx = static_cast<unsigned>(t.x); hx = static_cast<unsigned>(t.y); }
        else pk = gh16[row];      // raw (int8 g << 8 | int8 h): unpacked by unpack() where it is consumed, not here — the unpacking
                                  // instructions would wait for the load at the point of issue and stall the whole pipeline
      } else if (FLT) {
        const int4 e = reinterpret_cast<const int4*>(idx)[begin + p];
        row = e.x; gx = static_cast<unsigned>(e.y); hx = static_cast<unsigned>(e.z);
      } else {
        const int2 e = reinterpret_cast<const int2*>(idx)[begin + p];
        row = e.x; pk = static_cast<unsigned>(e.y);
      }
    };
    auto unpack = [](unsigned w) -> unsigned {      // -> (g << 16) + h, the shared-memory update word
      if (ROOT && !FLT) return (static_cast<unsigned>(static_cast<int>(static_cast<signed char>(w >> 8))) << 16) + (w & 0xffu);
      return w;
    };
    auto load_bins = [&](int row_l, int p0, int unit, unsigned (&bb)[U]) {
      if (ROOT) {
        // rows of a group are consecutive: one pointer per group, constant offsets per load
        const uint8_t* gp = rmc + static_cast<size_t>(static_cast<unsigned>(begin + p0 + r)) * rm_stride;
        #pragma unroll
        for (int u = 0; u < U; ++u) {
          const uint8_t* src = gp + static_cast<size_t>((unit * U + u) * R) * rm_stride;
          if (BPL == 4) bb[u] = __ldg(reinterpret_cast<const unsigned int*>(src));
          else bb[u] = __ldg(reinterpret_cast<const unsigned short*>(src));
        }
      } else {
        #pragma unroll
        for (int u = 0; u < U; ++u) {
          const unsigned row = static_cast<unsigned>(__shfl_sync(0xffffffffu, row_l, (unit * U + u) * R + r));
          const uint8_t* src = rmc + static_cast<size_t>(row) * rm_stride;
          if (BPL == 4) bb[u] = __ldg(reinterpret_cast<const unsigned int*>(src));
          else bb[u] = __ldg(reinterpret_cast<const unsigned short*>(src));
        }
      }
    };
    if (ng > 0) {
      int row_cur = 0, row_nxt = 0;
      unsigned pk_cur = 0, gx_cur = 0, hx_cur = 0, pk_nxt = 0, gx_nxt = 0, hx_nxt = 0;
      load_entry(first + lane, row_cur, pk_cur, gx_cur, hx_cur);
      if (ng > 1) load_entry(first + gstride + lane, row_nxt, pk_nxt, gx_nxt, hx_nxt);
      pk_cur = unpack(pk_cur);
      unsigned b_cur[U];
      load_bins(row_cur, first, 0, b_cur);
      #pragma unroll 1
      for (int j = 0; j < ng; ++j) {
        const int p0 = first + j * gstride;
        int row_nn = 0;
        unsigned pk_nn = 0, gx_nn = 0, hx_nn = 0;
        #pragma unroll
        for (int unit = 0; unit < NU; ++unit) {
          unsigned b_nxt[U];
          if (unit + 1 < NU) {
            load_bins(row_cur, p0, unit + 1, b_nxt);
          } else if (j + 1 < ng) {
            load_bins(row_nxt, p0 + gstride, 0, b_nxt);
            if (j + 2 < ng) load_entry(p0 + 2 * gstride + lane, row_nn, pk_nn, gx_nn, hx_nn);
          } else {
            #pragma unroll
            for (int u = 0; u < U; ++u) b_nxt[u] = 0;
          }
          #pragma unroll
          for (int u = 0; u < U; ++u) {
            const int srcl = (unit * U + u) * R + r;
            unsigned pk = 0, gx = 0, hx = 0;
            if (FLT) { gx = __shfl_sync(0xffffffffu, gx_cur, srcl); hx = __shfl_sync(0xffffffffu, hx_cur, srcl); }
            else pk = __shfl_sync(0xffffffffu, pk_cur, srcl);
            HistUpdate<BPL, FLT>(b_cur[u], lane_addr, pk, gx, hx);
          }
          #pragma unroll
          for (int u = 0; u < U; ++u) b_cur[u] = b_nxt[u];
        }
        row_cur = row_nxt; row_nxt = row_nn;
        pk_cur = unpack(pk_nxt); gx_cur = gx_nxt; hx_cur = hx_nxt;      // loaded one group ago
        pk_nxt = pk_nn; gx_nxt = gx_nn; hx_nxt = hx_nn;
      }
    }
    // ---- the warp's partial last group, if any: per-row predicates
    const int ptail = first + ng * gstride;
    if (ptail < sub_hi) {
      const int p = ptail + lane;
      const bool v = p < sub_hi;
      int row_l = -1;
      unsigned pk_l = 0, gx_l = 0, hx_l = 0;
      if (v) { load_entry(p, row_l, pk_l, gx_l, hx_l); pk_l = unpack(pk_l); }
      #pragma unroll 1
      for (int s = 0; s < STEPS; ++s) {
        const int srcl = s * R + r;
        const int row = __shfl_sync(0xffffffffu, row_l, srcl);
        unsigned pk = 0, gx = 0, hx = 0;
        if (FLT) { gx = __shfl_sync(0xffffffffu, gx_l, srcl); hx = __shfl_sync(0xffffffffu, hx_l, srcl); }
        else pk = __shfl_sync(0xffffffffu, pk_l, srcl);
        if (row >= 0 && kmask) {
          const uint8_t* src = rmc + static_cast<size_t>(static_cast<unsigned>(row)) * rm_stride;
          unsigned bb;
          if (BPL == 4) bb = __ldg(reinterpret_cast<const unsigned int*>(src));
          else bb = __ldg(reinterpret_cast<const unsigned short*>(src));
          HistUpdate<BPL, FLT>(bb, lane_addr, pk, gx, hx);
        }
      }
    }
    __syncthreads();
    if (t_flush != nullptr && threadIdx.x == 0) asm volatile("mov.u64 %0, %globaltimer;" : "=l"(*t_flush));      // timeline experiments
    HistFlush<R, BPL, FLT, PACK>(hsm, work, F, tile);
    __syncthreads();
  }
}

template <int R, int BPL, bool FLT, bool ROOT>
__global__ void __launch_bounds__(kHistThreads, 1)
k_hist(const uint8_t* __restrict__ rm, unsigned rm_stride, const uint16_t* __restrict__ gh16, const int2* __restrict__ ghfix,
       const int* __restrict__ idx0, const int* __restrict__ idx1, const LeafDev* __restrict__ leaves,
       TreeCtrl* __restrict__ ctrl, long long* __restrict__ work, int F, int rows_per_flush) {
  extern __shared__ unsigned int hsm_raw[];
  constexpr int LPR = 32 / R;                 // lanes per row
  constexpr int FT = LPR * BPL;               // features per tile
  constexpr int NARR = FLT ? 2 * BPL : BPL;   // 32 KB arrays (always 4)
  static_assert(NARR == 4, "four 32 KB arrays");
  PdlLaunchDependents();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  const int q = lane % LPR, r = lane / LPR;
  const int tile = blockIdx.y;
  // 32 KB-aligned window inside the dynamic shared memory
  const unsigned raw_addr = static_cast<unsigned>(__cvta_generic_to_shared(hsm_raw));
  const unsigned sbase = (raw_addr + kHistArrayBytes - 1) & ~(kHistArrayBytes - 1);
  unsigned int* hsm = hsm_raw + ((sbase - raw_addr) >> 2);
  const unsigned lane_addr = sbase + lane * 4;
  // A lane works if its first feature byte is a real feature (when F is not a multiple of BPL the last working lane also
  // updates its zero-padded bytes, bin 0 of features >= F, which the flush ignores).  In the hot loop the other lanes are
  // not predicated off: they read the tile's first feature bytes instead and update their own columns, which the flush
  // ignores as well — no branch, no hot address.
  const bool kmask = (tile * FT + q * BPL) < F;
  const uint8_t* rmc = rm + static_cast<size_t>(tile) * FT + (kmask ? q : 0) * BPL;
  asm volatile("" : "+l"(rmc));      // keep the pointer in registers (no re-derivation from the parameter bank per load)
  // clearing the shared histograms overlaps the predecessor kernel's tail
  for (int i = threadIdx.x; i < NARR * kBins * 32; i += blockDim.x) hsm[i] = 0u;
  PdlWait();
  if (ctrl->done) return;
  if (!ROOT && !ctrl->find) return;
  const LeafDev& lf = leaves[ROOT ? 0 : ctrl->smaller];
  const int begin = lf.begin, count = lf.count, buf = lf.buf;
  if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) {
    atomicAdd(reinterpret_cast<unsigned long long*>(&ctrl->hist_rows), static_cast<unsigned long long>(count));
    atomicAdd(reinterpret_cast<unsigned long long*>(&ctrl->hist_launches), 1ULL);
  }
  int rows_per_cta = (count + gridDim.x - 1) / gridDim.x;
  rows_per_cta = max((rows_per_cta + 31) & ~31, kHistMinRowsPerCta);
  const int lo = blockIdx.x * rows_per_cta;
  const int hi = min(count, lo + rows_per_cta);
  if (lo >= hi) return;
  const int* idx = ROOT ? nullptr : (buf == 0 ? idx0 : idx1);
  __syncthreads();
  HistRows<R, BPL, FLT, ROOT>(rmc, rm_stride, gh16, ghfix, idx, begin, lo, hi, lane_addr, hsm, work, F, tile, rows_per_flush, kmask);
}

}  // namespace lgb
