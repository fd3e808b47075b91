// K1 (bin-boundary search on the sampled rows) and K2 (row binning) kernels + the Dataset host object.
// Algorithm specification: upstream BinMapper::FindBin / FindBinWithZeroAsOneBin / GreedyFindBin / ValueToBin and
// DatasetLoader::ConstructFromSampleData.  Reference call sites are listed in dataset.h.
#include "dataset.h"

#include <cub/device/device_segmented_radix_sort.cuh>

#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits>
#include <map>
#include <mutex>

namespace lgb {

int g_verbosity = 1;

int ResolveDevice(const Config& cfg) {
  auto it = cfg.raw.find("gpu_device_id");
  if (it != cfg.raw.end()) {
    int d = atoi(it->second.c_str());
    if (d >= 0) return d;
  }
  // one process per GPU under torchrun: LOCAL_RANK names this process's device
  const char* lr = getenv("LOCAL_RANK");
  int ndev = 0;
  LGB_CUDA(cudaGetDeviceCount(&ndev));
  LGB_CHECK(ndev > 0, "No CUDA device is visible: this library has no CPU path");
  if (lr != nullptr) return atoi(lr) % ndev;
  return 0;
}

// =====================================================================================================================
// K1: bin-boundary search.
//   K1a  canonicalise NaNs + count them, then CUB segmented radix sort of every feature's sampled non-zero values
//   K1b/c one thread per feature: distinct values + counts (with the implicit zero count), zero-as-one-bin split, greedy
//        equal-frequency sweep with big-count values pinned, cnt_in_bin, trivial-feature filter, default/most-freq bin.
// The sweep is inherently sequential per feature; parallelism is F.  It runs once per fit on <= 200k x F values.
// =====================================================================================================================

struct K1Params {
  int max_bin, min_data_in_bin, filter_cnt, pre_filter, use_missing, zero_as_missing, total_sample_cnt;
  int bstride;   // doubles per feature in the bounds / scratch arrays (>= max_bin + 2)
};

struct K1Meta {
  int num_bin, missing_type, is_trivial, default_bin, most_freq_bin, bin_type;
  double sparse_rate, min_val, max_val;
};

__global__ void k1_canon_nan(double* vals, const long long* seg_off, int* na_cnt, int nfeat) {
  const int f = blockIdx.y;
  if (f >= nfeat) return;
  const long long b = seg_off[f], e = seg_off[f + 1];
  int local = 0;
  for (long long i = b + blockIdx.x * (long long)blockDim.x + threadIdx.x; i < e; i += (long long)gridDim.x * blockDim.x) {
    double v = vals[i];
    if (v != v) { vals[i] = __longlong_as_double(0x7ff8000000000000LL); ++local; }   // +qNaN sorts after +inf in radix order
  }
  for (int o = 16; o > 0; o >>= 1) local += __shfl_down_sync(0xffffffffu, local, o);
  if ((threadIdx.x & 31) == 0 && local) atomicAdd(&na_cnt[f], local);
}

// A dense row-major sample [rows][nfeat] instead of the host-built column arrays: per column, how many values are not zero (NaN
// counts, -0.0 does not: SampledData.scala:70-77), then the values themselves into the column's segment.  The order inside a
// segment is whatever the atomics give; the segments are sorted next, so the result does not depend on it.
__global__ void k1_dense_count(const double* __restrict__ rows, long long n_elems, int nfeat, int* __restrict__ counts) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n_elems; i += (long long)gridDim.x * blockDim.x)
    if (rows[i] != 0.0) atomicAdd(&counts[static_cast<int>(i % nfeat)], 1);
}

__global__ void k1_dense_scatter(const double* __restrict__ rows, long long n_elems, int nfeat, const long long* __restrict__ seg_off,
                                 int* __restrict__ cursor, double* __restrict__ vals) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n_elems; i += (long long)gridDim.x * blockDim.x) {
    const double v = rows[i];
    if (v != 0.0) {
      const int f = static_cast<int>(i % nfeat);
      vals[seg_off[f] + atomicAdd(&cursor[f], 1)] = v;
    }
  }
}

__device__ __forceinline__ double UpperBoundOf(double a) { return nextafter(a, __longlong_as_double(0x7ff0000000000000LL)); }
__device__ __forceinline__ bool EqualOrdered(double a, double b) { return b <= UpperBoundOf(a); }
#define K1_ZERO_THRESHOLD ((double)1e-35f)
#define K1_INF __longlong_as_double(0x7ff0000000000000LL)

// Greedy equal-frequency bins over distinct values dv[0..nd) with counts cnt[]; writes bounds to out, returns how many.
__device__ int K1Greedy(const double* dv, const int* cnt, int nd, int max_bin, long long total_cnt, int min_data_in_bin,
                        double* out, double* upper, double* lower) {
  int nb = 0;
  if (nd <= max_bin) {
    int cur = 0;
    for (int i = 0; i < nd - 1; ++i) {
      cur += cnt[i];
      if (cur >= min_data_in_bin) {
        double val = UpperBoundOf((dv[i] + dv[i + 1]) / 2.0);
        if (nb == 0 || !EqualOrdered(out[nb - 1], val)) { out[nb++] = val; cur = 0; }
      }
    }
    out[nb++] = K1_INF;
    return nb;
  }
  if (min_data_in_bin > 0) {
    int cap = (int)(total_cnt / min_data_in_bin);
    if (cap < max_bin) max_bin = cap;
    if (max_bin < 1) max_bin = 1;
  }
  const double mean0 = (double)total_cnt / max_bin;
  int rest_bin_cnt = max_bin;
  int rest_sample_cnt = (int)total_cnt;
  for (int i = 0; i < nd; ++i) {
    if (cnt[i] >= mean0) { --rest_bin_cnt; rest_sample_cnt -= cnt[i]; }
  }
  double mean_bin_size = (double)rest_sample_cnt / rest_bin_cnt;
  for (int i = 0; i < max_bin; ++i) { upper[i] = K1_INF; lower[i] = K1_INF; }
  int bin_cnt = 0;
  lower[0] = dv[0];
  int cur = 0;
  for (int i = 0; i < nd - 1; ++i) {
    const bool big_i = cnt[i] >= mean0;
    const bool big_n = cnt[i + 1] >= mean0;
    if (!big_i) rest_sample_cnt -= cnt[i];
    cur += cnt[i];
    double half = mean_bin_size * 0.5f;
    if (half < 1.0) half = 1.0;
    if (big_i || cur >= mean_bin_size || (big_n && cur >= half)) {
      upper[bin_cnt] = dv[i];
      ++bin_cnt;
      lower[bin_cnt] = dv[i + 1];
      if (bin_cnt >= max_bin - 1) break;
      cur = 0;
      if (!big_i) { --rest_bin_cnt; mean_bin_size = rest_sample_cnt / (double)rest_bin_cnt; }
    }
  }
  ++bin_cnt;
  for (int i = 0; i < bin_cnt - 1; ++i) {
    double val = UpperBoundOf((upper[i] + lower[i + 1]) / 2.0);
    if (nb == 0 || !EqualOrdered(out[nb - 1], val)) out[nb++] = val;
  }
  out[nb++] = K1_INF;
  return nb;
}

__device__ int K1ZeroAsOneBin(const double* dv, const int* cnt, int nd, int max_bin, long long total_sample_cnt,
                              int min_data_in_bin, double* out, double* tmp, double* upper, double* lower) {
  int left_cnt_data = 0, cnt_zero = 0, right_cnt_data = 0;
  for (int i = 0; i < nd; ++i) {
    if (dv[i] <= -K1_ZERO_THRESHOLD) left_cnt_data += cnt[i];
    else if (dv[i] > K1_ZERO_THRESHOLD) right_cnt_data += cnt[i];
    else cnt_zero += cnt[i];
  }
  int left_cnt = -1;
  for (int i = 0; i < nd; ++i) if (dv[i] > -K1_ZERO_THRESHOLD) { left_cnt = i; break; }
  if (left_cnt < 0) left_cnt = nd;
  int nb = 0;
  if (left_cnt > 0 && max_bin > 1) {
    int left_max_bin = (int)((double)left_cnt_data / (double)(total_sample_cnt - cnt_zero) * (max_bin - 1));
    if (left_max_bin < 1) left_max_bin = 1;
    nb = K1Greedy(dv, cnt, left_cnt, left_max_bin, left_cnt_data, min_data_in_bin, out, upper, lower);
    if (nb > 0) out[nb - 1] = -K1_ZERO_THRESHOLD;
  }
  int right_start = -1;
  for (int i = left_cnt; i < nd; ++i) if (dv[i] > K1_ZERO_THRESHOLD) { right_start = i; break; }
  const int right_max_bin = max_bin - 1 - nb;
  if (right_start >= 0 && right_max_bin > 0) {
    int nr = K1Greedy(dv + right_start, cnt + right_start, nd - right_start, right_max_bin, right_cnt_data, min_data_in_bin,
                      tmp, upper, lower);
    out[nb++] = K1_ZERO_THRESHOLD;
    for (int i = 0; i < nr; ++i) out[nb++] = tmp[i];
  } else {
    out[nb++] = K1_INF;
  }
  return nb;
}

__device__ int K1ValueToBin(double value, const double* bounds, int num_bin, int missing_type) {
  if (value != value) {
    if (missing_type == kMissingNaN) return num_bin - 1;
    value = 0.0;
  }
  int l = 0, r = num_bin - 1;
  if (missing_type == kMissingNaN) r -= 1;
  while (l < r) {
    int m = (r + l - 1) / 2;
    if (value <= bounds[m]) r = m; else l = m + 1;
  }
  return l;
}

// ---- block-cooperative versions (kK1Threads threads): the reductions over the distinct values run in parallel; only the
// greedy sweep itself is sequential, on thread 0, over chunks staged in shared memory (global memory latency per element made
// the single-thread version take 74 ms for 200k distinct values).  Results are identical to the sequential functions above.
constexpr int kK1Threads = 256;
constexpr int kK1Chunk = 2048;
struct K1Shared {
  double dv[kK1Chunk + 1];
  int cnt[kK1Chunk + 1];
  long long red[kK1Threads / 32];
  long long bcast[4];
};
__device__ __forceinline__ long long K1BlockSum(long long v, K1Shared* sh) {
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) sh->red[threadIdx.x >> 5] = v;
  __syncthreads();
  long long t = 0;
  for (int w = 0; w < kK1Threads / 32; ++w) t += sh->red[w];
  return t;
}

// K1Greedy, cooperative.  Every thread returns the number of bounds written to out.
__device__ int K1GreedyBlock(const double* dv, const int* cnt, int nd, int max_bin, long long total_cnt, int min_data_in_bin,
                             double* out, double* upper, double* lower, K1Shared* sh) {
  const int tid = threadIdx.x;
  if (nd <= max_bin) {      // at most max_bin iterations: thread 0 alone
    if (tid == 0) sh->bcast[0] = K1Greedy(dv, cnt, nd, max_bin, total_cnt, min_data_in_bin, out, upper, lower);
    __syncthreads();
    const int nb = static_cast<int>(sh->bcast[0]);
    __syncthreads();
    return nb;
  }
  if (min_data_in_bin > 0) {
    int cap = (int)(total_cnt / min_data_in_bin);
    if (cap < max_bin) max_bin = cap;
    if (max_bin < 1) max_bin = 1;
  }
  const double mean0 = (double)total_cnt / max_bin;
  long long nbig = 0, sbig = 0;
  for (int i = tid; i < nd; i += kK1Threads)
    if (cnt[i] >= mean0) { ++nbig; sbig += cnt[i]; }
  nbig = K1BlockSum(nbig, sh);
  sbig = K1BlockSum(sbig, sh);
  for (int i = tid; i < max_bin; i += kK1Threads) { upper[i] = K1_INF; lower[i] = K1_INF; }
  __syncthreads();
  // the sweep: thread 0 keeps the state in registers; all threads stage the next chunk (plus one look-ahead element)
  int rest_bin_cnt = max_bin - static_cast<int>(nbig);
  int rest_sample_cnt = (int)total_cnt - static_cast<int>(sbig);
  double mean_bin_size = (double)rest_sample_cnt / rest_bin_cnt;
  int bin_cnt = 0, cur = 0;
  if (tid == 0) { lower[0] = dv[0]; sh->bcast[1] = 0; }
  for (int c0 = 0; c0 < nd - 1; c0 += kK1Chunk) {
    const int m = min(kK1Chunk, nd - 1 - c0);       // elements c0 .. c0 + m - 1 are visited, c0 + m is looked at
    for (int i = tid; i <= m; i += kK1Threads) { sh->dv[i] = dv[c0 + i]; sh->cnt[i] = cnt[c0 + i]; }
    __syncthreads();
    if (tid == 0) {
      for (int j = 0; j < m; ++j) {
        const int ci = sh->cnt[j];
        const bool big_i = ci >= mean0;
        const bool big_n = sh->cnt[j + 1] >= mean0;
        if (!big_i) rest_sample_cnt -= ci;
        cur += ci;
        double half = mean_bin_size * 0.5f;
        if (half < 1.0) half = 1.0;
        if (big_i || cur >= mean_bin_size || (big_n && cur >= half)) {
          upper[bin_cnt] = sh->dv[j];
          ++bin_cnt;
          lower[bin_cnt] = sh->dv[j + 1];
          if (bin_cnt >= max_bin - 1) { sh->bcast[1] = 1; break; }
          cur = 0;
          if (!big_i) { --rest_bin_cnt; mean_bin_size = rest_sample_cnt / (double)rest_bin_cnt; }
        }
      }
    }
    __syncthreads();
    if (sh->bcast[1]) break;
  }
  if (tid == 0) {
    int nb = 0;
    ++bin_cnt;
    for (int i = 0; i < bin_cnt - 1; ++i) {
      double val = UpperBoundOf((upper[i] + lower[i + 1]) / 2.0);
      if (nb == 0 || !EqualOrdered(out[nb - 1], val)) out[nb++] = val;
    }
    out[nb++] = K1_INF;
    sh->bcast[0] = nb;
  }
  __syncthreads();
  const int nb = static_cast<int>(sh->bcast[0]);
  __syncthreads();
  return nb;
}

// K1ZeroAsOneBin, cooperative.  dv is ascending, so "first index with dv > x" is the number of values <= x.
__device__ int K1ZeroAsOneBinBlock(const double* dv, const int* cnt, int nd, int max_bin, long long total_sample_cnt,
                                   int min_data_in_bin, double* out, double* tmp, double* upper, double* lower, K1Shared* sh) {
  const int tid = threadIdx.x;
  long long l_data = 0, z_data = 0, r_data = 0, n_le_neg = 0, n_le_pos = 0;
  for (int i = tid; i < nd; i += kK1Threads) {
    const double x = dv[i];
    const int c = cnt[i];
    if (x <= -K1_ZERO_THRESHOLD) l_data += c;
    else if (x > K1_ZERO_THRESHOLD) r_data += c;
    else z_data += c;
    if (!(x > -K1_ZERO_THRESHOLD)) ++n_le_neg;
    if (!(x > K1_ZERO_THRESHOLD)) ++n_le_pos;
  }
  const int left_cnt_data = static_cast<int>(K1BlockSum(l_data, sh));
  const int cnt_zero = static_cast<int>(K1BlockSum(z_data, sh));
  const int right_cnt_data = static_cast<int>(K1BlockSum(r_data, sh));
  const int left_cnt = static_cast<int>(K1BlockSum(n_le_neg, sh));                 // first i with dv[i] > -threshold (nd if none)
  const int first_pos = static_cast<int>(K1BlockSum(n_le_pos, sh));
  const int right_start = first_pos < nd ? first_pos : -1;                         // first i with dv[i] > +threshold
  int nb = 0;
  if (left_cnt > 0 && max_bin > 1) {
    int left_max_bin = (int)((double)left_cnt_data / (double)(total_sample_cnt - cnt_zero) * (max_bin - 1));
    if (left_max_bin < 1) left_max_bin = 1;
    nb = K1GreedyBlock(dv, cnt, left_cnt, left_max_bin, left_cnt_data, min_data_in_bin, out, upper, lower, sh);
    if (nb > 0 && tid == 0) out[nb - 1] = -K1_ZERO_THRESHOLD;
    __syncthreads();
  }
  const int right_max_bin = max_bin - 1 - nb;
  if (right_start >= 0 && right_max_bin > 0) {
    const int nr = K1GreedyBlock(dv + right_start, cnt + right_start, nd - right_start, right_max_bin, right_cnt_data, min_data_in_bin,
                                 tmp, upper, lower, sh);
    if (tid == 0) {
      out[nb] = K1_ZERO_THRESHOLD;
      for (int i = 0; i < nr; ++i) out[nb + 1 + i] = tmp[i];
    }
    nb += 1 + nr;
  } else {
    if (tid == 0) out[nb] = K1_INF;
    nb += 1;
  }
  __syncthreads();
  return nb;
}

// Categorical feature (BinMapper::FindBin, bin_type == CategoricalBin), by the whole CTA.  dv / dc: the distinct sampled values
// (ascending) and their counts; ci: scratch for the category values.  The values are cut to ints (negative ones, and values beyond
// int, count as missing), equal neighbours merged, and the categories taken by falling count (ties: the smaller value first — the
// stable sort upstream) until 99 % of the non-missing sample is covered AND max_bin bins exist; a category seen fewer than
// min_data_in_bin times ends the list (except the first two).  Bin 0 collects everything else.  Every pick is a block-wide
// arg-max over the categories not taken yet.  out[b]: category of bin b (out[0] = -1); c[b]: sample rows of bin b.  Returns
// num_bin (0: no non-missing value; kMaxBinsPerFeature + 1: more categories than the uint8 bin matrix holds).
__device__ int K1CategoricalBlock(const double* dv, int* dc, int* ci, int nd, int na_cnt, const K1Params& p, double* out, int* c,
                                  int* missing_type) {
  __shared__ int s_n, s_na;
  __shared__ long long s_best[kK1Threads / 32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) {
    int n = 0, na = na_cnt;
    for (int i = 0; i < nd; ++i) {
      const double x = dv[i];
      const int val = x < 2147483648.0 ? static_cast<int>(x) : -1;
      const int cnt = dc[i];
      if (val < 0) na += cnt;
      else if (n == 0 || val != ci[n - 1]) { ci[n] = val; dc[n] = cnt; ++n; }      // in place: n <= i
      else dc[n - 1] += cnt;
    }
    s_n = n; s_na = na;
  }
  __syncthreads();
  const int n = s_n, na = s_na;
  const int rest_cnt = p.total_sample_cnt - na;
  if (rest_cnt <= 0) return 0;
  const int cut_cnt = static_cast<int>(static_cast<double>(__fmul_rn(static_cast<float>(rest_cnt), 0.99f)) + 0.5f);
  int distinct_cnt = n;
  if (na > 0) ++distinct_cnt;
  const int max_bin = min(distinct_cnt, p.max_bin);
  int num_bin = 1, used_cnt = 0, cur_cat = 0;
  if (tid == 0) { out[0] = -1.0; c[0] = 0; }
  while (cur_cat < n && (used_cnt < cut_cnt || num_bin < max_bin)) {
    // arg-max of (count, then the smaller index) over the categories not taken yet (taken: count -1)
    long long best = -1;
    for (int i = tid; i < n; i += kK1Threads) {
      const int cnt = dc[i];
      if (cnt >= 0) { const long long key = (static_cast<long long>(cnt) << 32) | static_cast<unsigned>(0x7fffffff - i); if (key > best) best = key; }
    }
    for (int o = 16; o > 0; o >>= 1) { const long long t = __shfl_xor_sync(0xffffffffu, best, o); if (t > best) best = t; }
    if (lane == 0) s_best[warp] = best;
    __syncthreads();
    best = s_best[0];
    for (int w = 1; w < kK1Threads / 32; ++w) if (s_best[w] > best) best = s_best[w];
    __syncthreads();
    const int cnt = static_cast<int>(best >> 32), idx = 0x7fffffff - static_cast<int>(best & 0xffffffffLL);
    if (cnt < p.min_data_in_bin && cur_cat > 1) break;
    if (num_bin >= kMaxBinsPerFeature) return kMaxBinsPerFeature + 1;
    if (tid == 0) { out[num_bin] = static_cast<double>(ci[idx]); c[num_bin] = cnt; dc[idx] = -1; }
    __syncthreads();
    used_cnt += cnt;
    ++num_bin;
    ++cur_cat;
  }
  *missing_type = (cur_cat == n && na == 0) ? kMissingNone : kMissingNaN;
  if (tid == 0) c[0] = p.total_sample_cnt - used_cnt;
  __syncthreads();
  return num_bin;
}

// sorted: every feature's segment sorted ascending, NaNs last.  distinct/counts/starts: scratch with room for n_f + 2 per
// feature (offset seg_off[f] + 2 f).  bounds/tmp/upper/lower/cib: [F][bstride].
// One CTA per feature.  The O(n_sample) part — runs of (ulp-)equal values -> distinct values + counts, with the implicit
// zero entry inserted where the sorted values cross zero — is done by the whole CTA (flag, block scan with carry, gather);
// the greedy bin construction is inherently sequential over the distinct values and runs on thread 0 with register
// accumulators (no read-modify-write of global memory inside its loops).
__global__ void __launch_bounds__(kK1Threads) k1_find_bin(const double* sorted, const long long* seg_off, const int* na_cnt_arr, int nfeat,
                                                        K1Params p, double* distinct, int* counts, int* starts, double* bounds, double* tmp,
                                                        double* upper, double* lower, int* cib, K1Meta* meta,
                                                        const unsigned char* __restrict__ is_cat, const int* __restrict__ max_bin_by_feature) {
  const int f = blockIdx.x;
  if (f >= nfeat) return;
  if (max_bin_by_feature != nullptr) p.max_bin = max_bin_by_feature[f];      // p is this CTA's own copy
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  __shared__ int s_warp[kK1Threads / 32];
  __shared__ int s_carry, s_nd;
  const double* v = sorted + seg_off[f];
  const int n_all = (int)(seg_off[f + 1] - seg_off[f]);
  int na_cnt = na_cnt_arr[f];
  const int nv = n_all - na_cnt;       // non-NaN values
  int missing_type;
  if (!p.use_missing) missing_type = kMissingNone;
  else if (p.zero_as_missing) missing_type = kMissingZero;
  else if (na_cnt == 0) missing_type = kMissingNone;
  else missing_type = kMissingNaN;
  if (missing_type != kMissingNaN) na_cnt = 0;
  const int zero_cnt = p.total_sample_cnt - nv - na_cnt;
  double* dv = distinct + seg_off[f] + 2LL * f;
  int* dc = counts + seg_off[f] + 2LL * f;
  int* st = starts + seg_off[f] + 2LL * f;
  // ---- pass 1: entry index of every run start; st[g] = first position of the run (a zero entry gets -1)
  const bool lead_zero = nv == 0 || (v[0] > 0.0 && zero_cnt > 0);
  if (tid == 0) { s_carry = lead_zero ? 1 : 0; if (lead_zero) st[0] = -1; }
  __syncthreads();
  for (int c0 = 0; c0 < nv; c0 += kK1Threads) {
    const int i = c0 + tid;
    bool start = false, cross = false;
    if (i < nv) {
      if (i == 0) start = true;
      else {
        const double a = v[i - 1], b = v[i];
        start = !EqualOrdered(a, b);
        cross = start && a < 0.0 && b > 0.0;
      }
    }
    const int inc = (start ? 1 : 0) + (cross ? 1 : 0);
    int incl = inc;
    for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    int before = 0, total = 0;
    for (int w = 0; w < kK1Threads / 32; ++w) { if (w < warp) before += s_warp[w]; total += s_warp[w]; }
    const int carry = s_carry;
    if (start) {
      const int g = carry + before + incl - 1;
      st[g] = i;
      if (cross) st[g - 1] = -1;
    }
    __syncthreads();
    if (tid == 0) s_carry = carry + total;
    __syncthreads();
  }
  if (tid == 0) {
    int nd = s_carry;
    if (nv > 0 && v[nv - 1] < 0.0 && zero_cnt > 0) { st[nd] = -1; ++nd; }
    s_nd = nd;
  }
  __syncthreads();
  const int nd = s_nd;
  // ---- pass 2: a run's value is its last element (what the sequential sweep leaves behind), its count the run length
  for (int g = tid; g < nd; g += kK1Threads) {
    const int b0 = st[g];
    if (b0 < 0) { dv[g] = 0.0; dc[g] = zero_cnt; continue; }
    int e = nv;
    for (int h = g + 1; h < nd; ++h) { const int t = st[h]; if (t >= 0) { e = t; break; } }     // at most one zero entry in between
    dv[g] = v[e - 1];
    dc[g] = e - b0;
  }
  __syncthreads();
  __shared__ K1Shared k1sh;
  double* out = bounds + (long long)f * p.bstride;
  double* t = tmp + (long long)f * p.bstride;
  double* up = upper + (long long)f * p.bstride;
  double* lo = lower + (long long)f * p.bstride;
  int num_bin;
  const bool cat = is_cat != nullptr && is_cat[f] != 0;
  int* c = cib + (long long)f * p.bstride;
  if (cat) {
    num_bin = K1CategoricalBlock(dv, dc, st, nd, na_cnt, p, out, c, &missing_type);
  } else if (missing_type == kMissingZero) {
    num_bin = K1ZeroAsOneBinBlock(dv, dc, nd, p.max_bin, p.total_sample_cnt, p.min_data_in_bin, out, t, up, lo, &k1sh);
    if (num_bin == 2) missing_type = kMissingNone;
  } else if (missing_type == kMissingNone) {
    num_bin = K1ZeroAsOneBinBlock(dv, dc, nd, p.max_bin, p.total_sample_cnt, p.min_data_in_bin, out, t, up, lo, &k1sh);
  } else {
    num_bin = K1ZeroAsOneBinBlock(dv, dc, nd, p.max_bin - 1, p.total_sample_cnt - na_cnt, p.min_data_in_bin, out, t, up, lo, &k1sh);
    if (tid == 0) out[num_bin] = __longlong_as_double(0x7ff8000000000000LL);
    ++num_bin;
  }
  // rows of the sample per bin: every distinct value goes to the first bin whose upper bound is not below it (the last bin
  // otherwise) — what the sequential sweep computes, done with a binary search per value and integer atomics
  if (!cat) {
    for (int i = tid; i < num_bin; i += kK1Threads) c[i] = 0;
    __syncthreads();
    const int last = num_bin - 1;
    for (int i = tid; i < nd; i += kK1Threads) {
      const double x = dv[i];
      int l = 0, r = last;           // first j in [0, last) with !(x > out[j]), else last
      while (l < r) {
        const int mid = (l + r) >> 1;
        if (x > out[mid]) l = mid + 1; else r = mid;
      }
      atomicAdd(&c[l], dc[i]);
    }
  }
  __syncthreads();
  if (tid != 0) return;
  if (!cat && missing_type == kMissingNaN) c[num_bin - 1] = na_cnt;
  K1Meta m;
  m.min_val = dv[0];
  m.max_val = dv[nd - 1];
  m.bin_type = cat ? 1 : 0;
  int is_trivial = num_bin <= 1;
  const bool cat_overflow = cat && num_bin > kMaxBinsPerFeature;      // refused by the host
  if (!is_trivial && p.pre_filter && !cat_overflow) {
    // NeedFilter: numerical — no cut leaves filter_cnt sample rows on both sides; categorical — with at most two bins, no bin
    // against the rest does (more bins are never filtered)
    bool need = true;
    int sum_left = 0;
    if (!cat) {
      for (int i = 0; i + 1 < num_bin; ++i) {
        sum_left += c[i];
        if (sum_left >= p.filter_cnt && p.total_sample_cnt - sum_left >= p.filter_cnt) { need = false; break; }
      }
    } else if (num_bin <= 2) {
      for (int i = 0; i + 1 < num_bin; ++i) {
        sum_left = c[i];
        if (sum_left >= p.filter_cnt && p.total_sample_cnt - sum_left >= p.filter_cnt) { need = false; break; }
      }
    } else {
      need = false;
    }
    if (need) is_trivial = 1;
  }
  m.num_bin = num_bin;
  m.missing_type = missing_type;
  m.is_trivial = is_trivial;
  m.default_bin = 0;
  m.most_freq_bin = 0;
  m.sparse_rate = 1.0;
  if (!is_trivial && !cat_overflow) {
    if (cat) { for (int i = 1; i < num_bin; ++i) if (out[i] == 0.0) m.default_bin = i; }      // the bin of category 0, if it has one
    else m.default_bin = K1ValueToBin(0.0, out, num_bin, missing_type);
    int arg = 0;
    for (int i = 1; i < num_bin; ++i) if (c[i] > c[arg]) arg = i;
    const double max_sparse_rate = (double)c[arg] / p.total_sample_cnt;
    if (arg != m.default_bin && max_sparse_rate < 0.7) arg = m.default_bin;
    m.most_freq_bin = arg;
    m.sparse_rate = (double)c[arg] / p.total_sample_cnt;
  }
  meta[f] = m;
}

// =====================================================================================================================
// K2: row binning.  One CTA owns a 32-feature tile and walks 128-row blocks (aligned to absolute row numbers so that
// the column-major stores are 16-byte aligned).  Lanes map to features, so a warp reads 32 consecutive doubles of one input
// row.  The tile's bin boundaries sit in shared memory as [bin][feature]: whatever bins the 32 lanes of a warp probe, lane l
// reads bank pair 2l — the 64-bit loads of a search step never conflict (a [feature][bin] layout with an odd stride
// serialised them 3-4 ways).  A thread bins 16 consecutive rows of its feature — 16 searches in flight, since a search is a
// chain of up to 8 dependent shared-memory loads — and stores them as ONE 128-bit word of the column-major matrix straight
// from registers; the row-major copy goes through a shared tile (conflict-free both ways) and leaves as 16 features of one
// row per 128-bit store.
// Algorithmic bytes: N * F * (8 read + 1 write) (+1 for the second layout).
// =====================================================================================================================

constexpr int K2_ROWS = 128;
constexpr int K2_FT = 32;

// CAT: the dataset has categorical or wide features.  A categorical feature's row of `bounds` holds its kept categories in
// ascending order and `catbin` the bin of each (BinMapper::ValueToBin, categorical: the value is cut to an int; NaN, negative and
// unlisted values are bin 0): the same lower-bound search finds the slot, an equality test and one byte load finish it.
// A column of a wide feature (more than 256 bins, dataset.h) searches the feature's bounds in global memory (16 halvings) and
// writes the byte that codes the logical bin for this column.
template <bool CAT>
__global__ void __launch_bounds__(256) k2_bin_rows(const double* __restrict__ data, int nrow, int ncol, long long start_row,
                                                   const double* __restrict__ bounds, const FeatureMetaDev* __restrict__ fmeta,
                                                   int f_used, uint8_t* __restrict__ cm, long long n_pad,
                                                   uint8_t* __restrict__ rm, int rm_stride, const uint8_t* __restrict__ catbin,
                                                   const double* __restrict__ wide_bounds) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double* sb = reinterpret_cast<double*>(smem_raw);                                       // [256 bins][32 features]
  uint8_t* tile_rm = smem_raw + sizeof(double) * K2_FT * kMaxBinsPerFeature;              // [128 rows][32 features]
  uint8_t* scb = tile_rm + K2_FT * K2_ROWS;                                               // CAT: [256 slots][32 features]
  const int ft = blockIdx.y;
  const int f0 = ft * K2_FT;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < K2_FT * kMaxBinsPerFeature; i += blockDim.x) {
    const int f = i >> 8, b = i & 255;      // consecutive threads read consecutive bins of one feature (coalesced)
    sb[b * K2_FT + f] = (f0 + f < f_used) ? bounds[(long long)(f0 + f) * kMaxBinsPerFeature + b] : 0.0;
    if (CAT) scb[b * K2_FT + f] = (f0 + f < f_used && catbin != nullptr) ? catbin[(long long)(f0 + f) * kMaxBinsPerFeature + b] : 0;
  }
  const int fi = f0 + lane;
  int num_bin = 1, missing_type = 0, real = 0;
  bool is_cat = false;
  if (fi < f_used) {
    num_bin = fmeta[fi].num_bin; missing_type = fmeta[fi].missing_type; real = fmeta[fi].real_index;
    if (CAT && fmeta[fi].bin_type != 0) { is_cat = true; num_bin -= 1; missing_type = kMissingNone; }      // search the num_bin - 1 categories
  }
  bool is_wide = false;
  int wide_c = 0, wide_nb = 0, wide_miss = -1, wide_mt = 0;
  const double* wb = nullptr;
  if (CAT && fi < f_used && fmeta[fi].wide_cols != 0) {
    const int wc = fmeta[fi].wide_cols;
    wide_c = wc > 0 ? 0 : -wc;
    const FeatureMetaDev first = fmeta[fi - wide_c];      // the feature's first column carries the logical description
    is_wide = true;
    wide_nb = first.num_bin; wide_mt = first.missing_type;
    wide_miss = wide_mt == kMissingNaN ? wide_nb - 1 : wide_mt == kMissingZero ? first.default_bin : -1;
    wb = wide_bounds + first.wide_off;
    num_bin = 1; missing_type = kMissingNone;      // the shared-memory search below does nothing for this lane
  }
  __syncthreads();
  const long long first_blk = start_row / K2_ROWS;
  const long long last_blk = (start_row + nrow - 1) / K2_ROWS;
  const double* myb = sb + lane;
  for (long long blk = first_blk + blockIdx.x; blk <= last_blk; blk += gridDim.x) {
    const long long r0 = blk * K2_ROWS;
    // 8 warps x 16 rows
    double val[16];
    int lo_[16], hi_[16];
    #pragma unroll
    for (int k = 0; k < 16; ++k) {
      const long long row = r0 + warp * 16 + k;
      const bool in = fi < f_used && row >= start_row && row < start_row + nrow;
      val[k] = in ? data[(row - start_row) * (long long)ncol + real] : 0.0;
      lo_[k] = 0;
      hi_[k] = in ? (missing_type == kMissingNaN ? num_bin - 2 : num_bin - 1) : 0;
    }
    unsigned int unlisted = 0;      // CAT: rows whose value is no category at all
    #pragma unroll
    for (int k = 0; k < 16; ++k) {
      if (CAT && is_cat) {
        const double v = val[k];
        const bool ok = v == v && v < 2147483648.0 && static_cast<int>(v) >= 0;
        if (!ok) unlisted |= 1u << k;
        val[k] = ok ? static_cast<double>(static_cast<int>(v)) : 0.0;
      } else if (val[k] != val[k]) {      // NaN: the last bin when the feature has a NaN bin, else it is binned as 0.0
        if (missing_type == kMissingNaN) { lo_[k] = hi_[k] = num_bin - 1; }
        else val[k] = 0.0;
      }
    }
    #pragma unroll
    for (int step = 0; step < 8; ++step) {      // lower bound over <= 256 bins: at most 8 halvings, branch-free and in lock step
      #pragma unroll
      for (int k = 0; k < 16; ++k) {
        const int l = lo_[k], r = hi_[k];
        const int m = (r + l - 1) >> 1;
        const bool go = l < r;
        const bool le = val[k] <= myb[(go ? m : 0) * K2_FT];
        hi_[k] = (go && le) ? m : r;
        lo_[k] = (go && !le) ? m + 1 : l;
      }
    }
    if (CAT && is_cat) {
      #pragma unroll
      for (int k = 0; k < 16; ++k) {
        const bool hit = !((unlisted >> k) & 1u) && myb[lo_[k] * K2_FT] == val[k];
        lo_[k] = hit ? static_cast<int>(scb[lo_[k] * K2_FT + lane]) : 0;
      }
    }
    if (CAT && is_wide) {      // BinMapper::ValueToBin over the feature's wide_nb bounds, then the byte of this column
      int wl[16], wr[16];
      #pragma unroll
      for (int k = 0; k < 16; ++k) {
        const long long row = r0 + warp * 16 + k;
        const bool in = row >= start_row && row < start_row + nrow;
        double v = in ? data[(row - start_row) * (long long)ncol + real] : 0.0;      // (val[k] was already cleaned of NaNs above)
        wl[k] = 0; wr[k] = wide_mt == kMissingNaN ? wide_nb - 2 : wide_nb - 1;
        if (v != v) {
          if (wide_mt == kMissingNaN) { wl[k] = wr[k] = wide_nb - 1; }
          else v = 0.0;
        }
        val[k] = v;
      }
      for (int step = 0; step < 16; ++step) {
        #pragma unroll
        for (int k = 0; k < 16; ++k) {
          const int l = wl[k], r = wr[k];
          const int m = (r + l - 1) >> 1;
          const bool go = l < r;
          const bool le = val[k] <= wb[go ? m : 0];
          wr[k] = (go && le) ? m : r;
          wl[k] = (go && !le) ? m + 1 : l;
        }
      }
      #pragma unroll
      for (int k = 0; k < 16; ++k) {
        const int logical = wl[k], home = logical / kWideLocalBins;
        lo_[k] = home == wide_c ? logical % kWideLocalBins + 1 : logical == wide_miss ? kWideMissing : home > wide_c ? kWideAbove : kWideBelow;
      }
    }
    unsigned int w[4];
    #pragma unroll
    for (int j = 0; j < 4; ++j)
      w[j] = static_cast<unsigned>(lo_[4 * j]) | (static_cast<unsigned>(lo_[4 * j + 1]) << 8) | (static_cast<unsigned>(lo_[4 * j + 2]) << 16) |
             (static_cast<unsigned>(lo_[4 * j + 3]) << 24);
    #pragma unroll
    for (int k = 0; k < 16; ++k) tile_rm[(warp * 16 + k) * K2_FT + lane] = static_cast<uint8_t>(lo_[k]);
    const bool full = (r0 >= start_row) && (r0 + K2_ROWS <= start_row + nrow);
    if (fi < f_used) {      // column-major: this thread's 16 rows of its feature
      uint8_t* dst = cm + (long long)fi * n_pad + r0 + warp * 16;
      if (full) {
        *reinterpret_cast<uint4*>(dst) = make_uint4(w[0], w[1], w[2], w[3]);
      } else {
        #pragma unroll
        for (int k = 0; k < 16; ++k) {
          const long long row = r0 + warp * 16 + k;
          if (row >= start_row && row < start_row + nrow) dst[k] = static_cast<uint8_t>(lo_[k]);
        }
      }
    }
    __syncthreads();
    {  // row-major: thread -> (row, 16-feature half)
      const int rr = threadIdx.x >> 1, half = threadIdx.x & 1;
      const long long row = r0 + rr;
      if (row >= start_row && row < start_row + nrow) {
        uint8_t* dst = rm + row * (long long)rm_stride + f0 + half * 16;
        *reinterpret_cast<uint4*>(dst) = *reinterpret_cast<const uint4*>(tile_rm + rr * K2_FT + half * 16);
      }
    }
    __syncthreads();
  }
}

// =====================================================================================================================
// Host side
// =====================================================================================================================

static BinLayout ChooseLayout(int f_used) {
  BinLayout L;
  if (f_used <= 32) { L.rows_per_step = 4; L.rm_stride = 32; }
  else if (f_used <= 64) { L.rows_per_step = 2; L.rm_stride = 64; }
  else { L.rows_per_step = 1; L.rm_stride = RoundUp(f_used, 128); }
  L.tile_features_q = 32 / L.rows_per_step * 4;
  L.tile_features_f = 32 / L.rows_per_step * 2;
  return L;
}

Dataset::~Dataset() { FreeSlots(); }

void Dataset::FinishSchema() {
  used.clear(); logical_used.clear(); logical_first_col.clear(); logical_num_cols.clear();
  real_to_inner.assign(num_total_features, -1);
  max_num_bin = 1;
  has_wide = false;
  std::vector<double> hwb;      // bounds of the wide features
  std::vector<int> wide_off(num_total_features, 0);
  for (int j = 0; j < num_total_features; ++j) {
    const BinMapperH& m = mappers[j];
    if (m.is_trivial) continue;
    LGB_CHECK(m.num_bin <= kMaxWideBins, "Feature %d has %d bins; at most %d are supported", j, m.num_bin, kMaxWideBins);
    LGB_CHECK(m.bin_type == 0 || m.num_bin <= kMaxBinsPerFeature, "Categorical feature %d has %d bins; at most %d are supported", j, m.num_bin,
              kMaxBinsPerFeature);
    const int cols = m.num_bin <= kMaxBinsPerFeature ? 1 : CeilDiv(m.num_bin, kWideLocalBins);
    real_to_inner[j] = static_cast<int>(used.size());
    logical_used.push_back(j);
    logical_first_col.push_back(static_cast<int>(used.size()));
    logical_num_cols.push_back(cols);
    for (int c = 0; c < cols; ++c) used.push_back(j);
    if (cols > 1) {
      has_wide = true;
      wide_off[j] = static_cast<int>(hwb.size());
      hwb.insert(hwb.end(), m.upper.begin(), m.upper.end());
    }
    max_num_bin = std::max(max_num_bin, cols > 1 ? kMaxBinsPerFeature : m.num_bin);
  }
  if (feature_names.empty())
    for (int j = 0; j < num_total_features; ++j) feature_names.push_back("Column_" + std::to_string(j));
  const int F = num_features();
  layout = ChooseLayout(std::max(F, 1));
  std::vector<double> hb(static_cast<size_t>(std::max(F, 1)) * kMaxBinsPerFeature, std::numeric_limits<double>::infinity());
  std::vector<FeatureMetaDev> hm(std::max(F, 1));
  std::vector<int> hr(std::max(F, 1), 0);
  std::vector<uint8_t> hcb(hb.size(), 0);
  has_categorical = false;
  for (int i = 0; i < F; ++i) {
    const BinMapperH& m = mappers[used[i]];
    const int first = real_to_inner[used[i]];
    if (m.num_bin > kMaxBinsPerFeature) {      // a column of a wide feature: its bounds live in d_wide_bounds
      const int S = CeilDiv(m.num_bin, kWideLocalBins), c = i - first;
      // the bin whose rows follow default_left: the NaN bin, or the zero bin under zero_as_missing
      const int miss = m.missing_type == kMissingNaN ? m.num_bin - 1 : m.missing_type == kMissingZero ? m.default_bin : -1;
      const int nan_local = miss < 0 ? 256 : (miss / kWideLocalBins == c ? miss % kWideLocalBins + 1 : kWideMissing);
      hm[i] = FeatureMetaDev{c == 0 ? m.num_bin : kMaxBinsPerFeature, m.missing_type, m.default_bin, m.most_freq_bin, used[i], 0,
                             c == 0 ? S : -c, nan_local, wide_off[used[i]], 0};
      hr[i] = used[i];
      continue;
    }
    if (m.bin_type != 0) {      // categorical: the kept categories in ascending order and the bin of each (the row-binning kernel's search)
      std::vector<std::pair<int, int>> cb;
      for (int b = 1; b < m.num_bin; ++b) cb.emplace_back(m.bin_2_cat[b], b);
      std::sort(cb.begin(), cb.end());
      for (size_t k = 0; k < cb.size(); ++k) {
        hb[static_cast<size_t>(i) * kMaxBinsPerFeature + k] = static_cast<double>(cb[k].first);
        hcb[static_cast<size_t>(i) * kMaxBinsPerFeature + k] = static_cast<uint8_t>(cb[k].second);
      }
      has_categorical = true;
    } else {
      for (int b = 0; b < m.num_bin; ++b) hb[static_cast<size_t>(i) * kMaxBinsPerFeature + b] = m.upper[b];
    }
    hm[i] = FeatureMetaDev{m.num_bin, m.missing_type, m.default_bin, m.most_freq_bin, used[i], m.bin_type, 0, 0, 0, 0};
    hr[i] = used[i];
  }
  LGB_CUDA(cudaSetDevice(device));
  d_bounds.Alloc(hb.size());
  LGB_CUDA(cudaMemcpy(d_bounds.p, hb.data(), hb.size() * sizeof(double), cudaMemcpyHostToDevice));
  d_fmeta.Alloc(hm.size());
  LGB_CUDA(cudaMemcpy(d_fmeta.p, hm.data(), hm.size() * sizeof(FeatureMetaDev), cudaMemcpyHostToDevice));
  if (has_wide) {
    d_wide_bounds.Alloc(hwb.size());
    LGB_CUDA(cudaMemcpy(d_wide_bounds.p, hwb.data(), hwb.size() * sizeof(double), cudaMemcpyHostToDevice));
  }
  if (has_categorical) {
    d_catbin.Alloc(hcb.size());
    LGB_CUDA(cudaMemcpy(d_catbin.p, hcb.data(), hcb.size(), cudaMemcpyHostToDevice));
  }
  d_real_index.Alloc(hr.size());
  LGB_CUDA(cudaMemcpy(d_real_index.p, hr.data(), hr.size() * sizeof(int), cudaMemcpyHostToDevice));
}

void Dataset::AllocateRows(int64_t n) {
  LGB_CHECK(n >= 0 && n < (int64_t(1) << 31) - 256, "num_data %lld out of range", (long long)n);
  num_data = n;
  n_pad = RoundUp<int64_t>(std::max<int64_t>(n, 1), 128);
  LGB_CUDA(cudaSetDevice(device));
  const int F = std::max(num_features(), 1);
  d_cm.AllocZero(static_cast<size_t>(F) * n_pad);
  d_rm.AllocZero(static_cast<size_t>(n_pad) * layout.rm_stride);
  label.assign(n, 0.0f);
  pushed_rows = 0;
  finished = false;
}

Dataset* Dataset::CreateFromSampledColumn(double** sample_data, int** sample_indices, int32_t ncol, const int* num_per_col,
                                          int32_t num_sample_row, int32_t num_local_row, int64_t num_dist_row,
                                          const char* params, const double* dense_rows) {
  (void)sample_indices;
  std::unique_ptr<Dataset> ds(new Dataset());
  ds->param_string = params ? params : "";
  ds->cfg.Set(ds->param_string);
  g_verbosity = ds->cfg.verbosity;
  const Config& c = ds->cfg;
  LGB_CHECK(ncol > 0, "ncol must be positive");
  LGB_CHECK(c.max_bin >= 2 && c.max_bin <= kMaxWideBins, "max_bin=%d is outside [2, %d]", c.max_bin, kMaxWideBins);
  std::vector<unsigned char> h_is_cat(ncol, 0);
  for (int j : c.categorical_feature) {
    LGB_CHECK(j >= 0 && j < ncol, "categorical_feature=%d is not a column index (0..%d)", j, ncol - 1);
    h_is_cat[j] = 1;
  }
  const bool any_cat = !c.categorical_feature.empty();
  ds->device = ResolveDevice(c);
  LGB_CUDA(cudaSetDevice(ds->device));
  ds->num_total_features = ncol;
  // ---- stage the column-major non-zero sample (SampledData.scala:30-40) into one device array; from a dense row-major sample the
  // device drops the zeros and transposes (the host then never builds the per-column arrays)
  DevBuf<double> d_rows;
  std::vector<int> dense_counts;
  const long long n_elems = static_cast<long long>(num_sample_row) * ncol;
  const unsigned dense_grid = static_cast<unsigned>(std::max<long long>(1, std::min<long long>(CeilDiv<long long>(n_elems, 256), 148 * 16)));
  if (dense_rows != nullptr) {
    LGB_CHECK(num_sample_row > 0, "num_sample_row must be positive");
    d_rows.Alloc(n_elems);
    LGB_CUDA(cudaMemcpy(d_rows.p, dense_rows, sizeof(double) * n_elems, cudaMemcpyHostToDevice));
    DevBuf<int> d_cnt;
    d_cnt.AllocZero(ncol);
    k1_dense_count<<<dense_grid, 256>>>(d_rows.p, n_elems, ncol, d_cnt.p);
    LGB_CUDA(cudaGetLastError());
    dense_counts.resize(ncol);
    LGB_CUDA(cudaMemcpy(dense_counts.data(), d_cnt.p, sizeof(int) * ncol, cudaMemcpyDeviceToHost));
    num_per_col = dense_counts.data();
  }
  std::vector<long long> off(ncol + 1, 0);
  for (int j = 0; j < ncol; ++j) {
    LGB_CHECK(num_per_col[j] >= 0 && num_per_col[j] <= num_sample_row, "num_per_col[%d]=%d out of range", j, num_per_col[j]);
    off[j + 1] = off[j] + num_per_col[j];
  }
  const long long total = off[ncol];
  DevBuf<double> d_vals, d_sorted, d_distinct, d_bnd, d_tmp, d_up, d_lo;
  DevBuf<long long> d_off;
  DevBuf<int> d_na, d_counts, d_starts, d_cib;
  DevBuf<K1Meta> d_meta;
  d_vals.Alloc(std::max<long long>(total, 1));
  d_sorted.Alloc(std::max<long long>(total, 1));
  d_off.Alloc(ncol + 1);
  LGB_CUDA(cudaMemcpy(d_off.p, off.data(), sizeof(long long) * (ncol + 1), cudaMemcpyHostToDevice));
  if (dense_rows != nullptr) {
    DevBuf<int> d_cursor;
    d_cursor.AllocZero(ncol);
    k1_dense_scatter<<<dense_grid, 256>>>(d_rows.p, n_elems, ncol, d_off.p, d_cursor.p, d_vals.p);
    LGB_CUDA(cudaGetLastError());
    LGB_CUDA(cudaDeviceSynchronize());
    d_rows.Free();
  } else {
    for (int j = 0; j < ncol; ++j)
      if (num_per_col[j] > 0)
        LGB_CUDA(cudaMemcpy(d_vals.p + off[j], sample_data[j], sizeof(double) * num_per_col[j], cudaMemcpyHostToDevice));
  }
  d_na.AllocZero(ncol);
  if (total > 0) {
    dim3 grid(static_cast<unsigned>(std::min<long long>(CeilDiv<long long>(num_sample_row, 256), 64)), ncol);
    k1_canon_nan<<<grid, 256>>>(d_vals.p, d_off.p, d_na.p, ncol);
    LGB_CUDA(cudaGetLastError());
    size_t tmp_bytes = 0;
    LGB_CUDA(cub::DeviceSegmentedRadixSort::SortKeys(nullptr, tmp_bytes, d_vals.p, d_sorted.p, total, ncol, d_off.p, d_off.p + 1));
    DevBuf<unsigned char> d_cubtmp;
    d_cubtmp.Alloc(std::max<size_t>(tmp_bytes, 16));
    LGB_CUDA(cub::DeviceSegmentedRadixSort::SortKeys(d_cubtmp.p, tmp_bytes, d_vals.p, d_sorted.p, total, ncol, d_off.p, d_off.p + 1));
  }
  K1Params p;
  p.max_bin = c.max_bin;
  p.min_data_in_bin = c.min_data_in_bin;
  // filter_cnt: the reference passes num_local_row = 1 and num_dist_row = N (ReferenceDatasetUtils.scala:44-45)
  const int64_t denom = num_dist_row > 0 ? num_dist_row : std::max<int64_t>(num_local_row, 1);
  p.filter_cnt = static_cast<int>(static_cast<double>(static_cast<size_t>(c.min_data_in_leaf) * static_cast<size_t>(num_sample_row)) / denom);
  p.pre_filter = c.feature_pre_filter;
  p.use_missing = c.use_missing;
  p.zero_as_missing = c.zero_as_missing;
  p.total_sample_cnt = num_sample_row;
  int widest = c.max_bin;
  DevBuf<int> d_mbf;
  if (!c.max_bin_by_feature.empty()) {      // DatasetLoader::ConstructFromSampleData: one max_bin per column
    LGB_CHECK(static_cast<int>(c.max_bin_by_feature.size()) == ncol, "max_bin_by_feature has %d values for %d columns",
              static_cast<int>(c.max_bin_by_feature.size()), ncol);
    for (int v : c.max_bin_by_feature) {
      LGB_CHECK(v >= 2 && v <= kMaxWideBins, "max_bin_by_feature value %d is outside [2, %d]", v, kMaxWideBins);
      widest = std::max(widest, v);
    }
    d_mbf.Alloc(ncol);
    LGB_CUDA(cudaMemcpy(d_mbf.p, c.max_bin_by_feature.data(), sizeof(int) * ncol, cudaMemcpyHostToDevice));
  }
  p.bstride = any_cat ? std::max(widest + 4, kMaxBinsPerFeature + 4) : widest + 4;      // a categorical feature may exceed max_bin bins
  DevBuf<unsigned char> d_is_cat;
  if (any_cat) {
    d_is_cat.Alloc(ncol);
    LGB_CUDA(cudaMemcpy(d_is_cat.p, h_is_cat.data(), ncol, cudaMemcpyHostToDevice));
  }
  d_distinct.Alloc(total + 2LL * ncol + 2);
  d_counts.Alloc(total + 2LL * ncol + 2);
  d_starts.Alloc(total + 2LL * ncol + 2);
  const size_t bs = static_cast<size_t>(ncol) * p.bstride;
  d_bnd.Alloc(bs); d_tmp.Alloc(bs); d_up.Alloc(bs); d_lo.Alloc(bs); d_cib.Alloc(bs);
  d_meta.Alloc(ncol);
  k1_find_bin<<<ncol, kK1Threads>>>(d_sorted.p, d_off.p, d_na.p, ncol, p, d_distinct.p, d_counts.p, d_starts.p, d_bnd.p, d_tmp.p,
                                    d_up.p, d_lo.p, d_cib.p, d_meta.p, any_cat ? d_is_cat.p : nullptr,
                                    c.max_bin_by_feature.empty() ? nullptr : d_mbf.p);
  LGB_CUDA(cudaGetLastError());
  LGB_CUDA(cudaDeviceSynchronize());
  std::vector<K1Meta> hmeta(ncol);
  std::vector<double> hb(bs);
  LGB_CUDA(cudaMemcpy(hmeta.data(), d_meta.p, sizeof(K1Meta) * ncol, cudaMemcpyDeviceToHost));
  LGB_CUDA(cudaMemcpy(hb.data(), d_bnd.p, sizeof(double) * bs, cudaMemcpyDeviceToHost));
  ds->mappers.resize(ncol);
  for (int j = 0; j < ncol; ++j) {
    BinMapperH& m = ds->mappers[j];
    m.num_bin = hmeta[j].num_bin; m.missing_type = hmeta[j].missing_type; m.is_trivial = hmeta[j].is_trivial;
    m.default_bin = hmeta[j].default_bin; m.most_freq_bin = hmeta[j].most_freq_bin; m.sparse_rate = hmeta[j].sparse_rate;
    m.min_val = hmeta[j].min_val; m.max_val = hmeta[j].max_val;
    m.bin_type = hmeta[j].bin_type;
    LGB_CHECK(m.bin_type == 0 || m.num_bin <= kMaxBinsPerFeature,
              "Categorical feature %d needs more than %d bins, which is not supported", j, kMaxBinsPerFeature);
    m.upper.assign(hb.begin() + static_cast<size_t>(j) * p.bstride, hb.begin() + static_cast<size_t>(j) * p.bstride + m.num_bin);
    if (m.bin_type != 0) {      // the "bounds" of a categorical feature are the categories of its bins
      m.bin_2_cat.resize(m.num_bin);
      for (int b = 0; b < m.num_bin; ++b) m.bin_2_cat[b] = static_cast<int>(m.upper[b]);
    }
  }
  ds->FinishSchema();
  ds->AllocateRows(num_local_row);
  return ds.release();
}

// ---- reference dataset serialisation: the byte[] the Scala side ships driver -> executors is opaque to it
// (ReferenceDatasetUtils.scala:99-138), so the format is ours: little-endian, versioned.
namespace {
struct Writer {
  std::string s;
  template <typename T> void Put(const T& v) { s.append(reinterpret_cast<const char*>(&v), sizeof(T)); }
  void PutStr(const std::string& v) { Put<int32_t>(static_cast<int32_t>(v.size())); s.append(v); }
};
struct Reader {
  const char* p; const char* e;
  template <typename T> T Get() { LGB_CHECK(p + sizeof(T) <= e, "Serialized reference dataset is truncated"); T v; memcpy(&v, p, sizeof(T)); p += sizeof(T); return v; }
  std::string GetStr() { int32_t n = Get<int32_t>(); LGB_CHECK(n >= 0 && p + n <= e, "Serialized reference dataset is truncated"); std::string v(p, p + n); p += n; return v; }
};
constexpr uint32_t kRefMagic = 0x4C42B200u;
}  // namespace

std::string Dataset::SerializeReference() const {
  Writer w;
  w.Put<uint32_t>(kRefMagic);
  w.Put<int32_t>(2);
  w.PutStr(param_string);
  w.Put<int32_t>(num_total_features);
  for (const auto& m : mappers) {
    w.Put<int32_t>(m.num_bin); w.Put<int32_t>(m.missing_type); w.Put<int32_t>(m.is_trivial); w.Put<int32_t>(m.default_bin);
    w.Put<int32_t>(m.most_freq_bin); w.Put<double>(m.sparse_rate); w.Put<double>(m.min_val); w.Put<double>(m.max_val);
    w.Put<int32_t>(m.bin_type);
    for (double b : m.upper) w.Put<double>(b);
  }
  for (const auto& n : feature_names) w.PutStr(n);
  return w.s;
}

Dataset* Dataset::CreateFromSerializedReference(const void* buf, int32_t len, int64_t num_row, const char* params) {
  Reader r{static_cast<const char*>(buf), static_cast<const char*>(buf) + len};
  LGB_CHECK(r.Get<uint32_t>() == kRefMagic, "Buffer is not a serialized reference dataset of this library");
  LGB_CHECK(r.Get<int32_t>() == 2, "Unsupported serialized reference dataset version");
  std::unique_ptr<Dataset> ds(new Dataset());
  ds->param_string = r.GetStr();
  ds->cfg.Set(ds->param_string);         // the parameters the reference was built with (driver side) ...
  ds->cfg.Set(params ? params : "");     // ... overridden by the ones given here (the executor's own: gpu_device_id, verbosity, ...)
  g_verbosity = ds->cfg.verbosity;
  ds->device = ResolveDevice(ds->cfg);
  ds->num_total_features = r.Get<int32_t>();
  LGB_CHECK(ds->num_total_features > 0 && ds->num_total_features < (1 << 24), "Corrupt serialized reference dataset");
  ds->mappers.resize(ds->num_total_features);
  for (auto& m : ds->mappers) {
    m.num_bin = r.Get<int32_t>(); m.missing_type = r.Get<int32_t>(); m.is_trivial = r.Get<int32_t>();
    m.default_bin = r.Get<int32_t>(); m.most_freq_bin = r.Get<int32_t>(); m.sparse_rate = r.Get<double>();
    m.min_val = r.Get<double>(); m.max_val = r.Get<double>();
    m.bin_type = r.Get<int32_t>();
    LGB_CHECK(m.num_bin >= 0 && m.num_bin <= 65536 && (m.bin_type == 0 || m.bin_type == 1), "Corrupt serialized reference dataset");
    m.upper.resize(m.num_bin);
    for (auto& b : m.upper) b = r.Get<double>();
    if (m.bin_type != 0) {
      m.bin_2_cat.resize(m.num_bin);
      for (int b = 0; b < m.num_bin; ++b) m.bin_2_cat[b] = static_cast<int>(m.upper[b]);
    }
  }
  ds->feature_names.resize(ds->num_total_features);
  for (auto& n : ds->feature_names) n = r.GetStr();
  ds->FinishSchema();
  ds->AllocateRows(num_row);
  return ds.release();
}

Dataset* Dataset::CreateByReference(int64_t num_row) const {
  std::unique_ptr<Dataset> ds(new Dataset());
  ds->param_string = param_string;
  ds->cfg = cfg;
  ds->device = device;
  ds->num_total_features = num_total_features;
  ds->mappers = mappers;
  ds->feature_names = feature_names;
  ds->FinishSchema();
  ds->AllocateRows(num_row);
  return ds.release();
}

void Dataset::SetFeatureNames(const char** names, int n) {
  LGB_CHECK(n == num_total_features, "Expected %d feature names, got %d", num_total_features, n);
  feature_names.assign(names, names + n);
}

constexpr size_t kStageBytes = size_t(4) << 20;

// Page-locked staging buffers are expensive to create (tens of milliseconds per 64 MB), so they are kept for the life of the
// process and handed from one dataset to the next (a fit creates a dataset per executor; repeated fits reuse the buffers).
namespace {
struct PinnedCache {
  std::mutex mu;
  std::multimap<size_t, void*> free_list;
  size_t cached_bytes = 0;
  static constexpr size_t kMaxCached = size_t(1) << 30;
  void* Get(size_t bytes) {
    {
      std::lock_guard<std::mutex> g(mu);
      auto it = free_list.find(bytes);
      if (it != free_list.end()) { void* p = it->second; free_list.erase(it); cached_bytes -= bytes; return p; }
    }
    void* p = nullptr;
    LGB_CUDA(cudaMallocHost(&p, bytes));
    return p;
  }
  void Put(void* p, size_t bytes) {
    {
      std::lock_guard<std::mutex> g(mu);
      if (cached_bytes + bytes <= kMaxCached) { free_list.emplace(bytes, p); cached_bytes += bytes; return; }
    }
    cudaFreeHost(p);
  }
};
PinnedCache& Pinned() { static PinnedCache* c = new PinnedCache(); return *c; }   // never destroyed: outlives the CUDA context teardown order
}  // namespace

void Dataset::EnsureSlots(int n) {
  std::lock_guard<std::mutex> g(mu_);
  LGB_CUDA(cudaSetDevice(device));
  while (static_cast<int>(slots_.size()) < n) {
    PushSlot s;
    s.cap_rows = std::max<size_t>(kStageBytes / (sizeof(double) * num_total_features), 1);
    const size_t bytes = s.cap_rows * num_total_features * sizeof(double);
    for (int k = 0; k < 2; ++k) {
      s.pinned[k] = static_cast<double*>(Pinned().Get(bytes));
      s.pinned_bytes = bytes;
      LGB_CUDA(cudaMalloc(&s.dstage[k], bytes));
      LGB_CUDA(cudaEventCreateWithFlags(&s.ev[k], cudaEventDisableTiming));
    }
    LGB_CUDA(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
    slots_.push_back(s);
  }
}

void Dataset::FreeSlots() {
  cudaSetDevice(device);
  for (auto& s : slots_) {
    if (s.stream) cudaStreamSynchronize(s.stream);
    for (int k = 0; k < 2; ++k) {
      if (s.pinned[k]) Pinned().Put(s.pinned[k], s.pinned_bytes);
      if (s.dstage[k]) cudaFree(s.dstage[k]);
      if (s.ev[k]) cudaEventDestroy(s.ev[k]);
    }
    if (s.stream) cudaStreamDestroy(s.stream);
  }
  slots_.clear();
}

void Dataset::InitStreaming(bool has_w, bool has_init, bool has_q, int nclasses, int nthreads) {
  LGB_CHECK(nclasses >= 1, "InitStreaming: nclasses must be positive");
  num_init_score_classes = nclasses;
  has_weight = has_w; has_init_score = has_init; has_query = has_q;
  if (has_w) weight.assign(num_data, 0.0f);
  if (has_init) init_score.assign(num_data * nclasses, 0.0);      // class-major [nclasses][num_data], upstream Metadata's layout
  if (has_q) query_id.assign(num_data, 0);
  EnsureSlots(std::max(nthreads, 1));
  streaming_inited = true;
}

// One launch of the row-binning kernel over `nrow` staged rows that start at absolute row `start_row`.
void Dataset::LaunchBinRows(const double* staged, int nrow, long long start_row, cudaStream_t stream) {
  const int F = num_features();
  const long long nblk = (start_row + nrow - 1) / K2_ROWS - start_row / K2_ROWS + 1;
  const int ftiles = CeilDiv(F, K2_FT);
  dim3 grid(static_cast<unsigned>(std::min<long long>(nblk, std::max(1, 148 * 4 / ftiles))), ftiles);
  const bool ext = has_categorical || has_wide;
  const size_t smem = sizeof(double) * K2_FT * kMaxBinsPerFeature + K2_FT * K2_ROWS + (ext ? K2_FT * kMaxBinsPerFeature : 0);
  static bool attr_set[2] = {false, false};
  if (!attr_set[ext ? 1 : 0]) {
    if (ext) LGB_CUDA(cudaFuncSetAttribute(k2_bin_rows<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    else LGB_CUDA(cudaFuncSetAttribute(k2_bin_rows<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
    attr_set[ext ? 1 : 0] = true;
  }
  if (ext)
    k2_bin_rows<true><<<grid, 256, smem, stream>>>(staged, nrow, num_total_features, start_row, d_bounds.p, d_fmeta.p, F, d_cm.p, n_pad, d_rm.p,
                                                   layout.rm_stride, has_categorical ? d_catbin.p : nullptr, has_wide ? d_wide_bounds.p : nullptr);
  else
    k2_bin_rows<false><<<grid, 256, smem, stream>>>(staged, nrow, num_total_features, start_row, d_bounds.p, d_fmeta.p, F, d_cm.p, n_pad, d_rm.p,
                                                    layout.rm_stride, nullptr, nullptr);
}

void Dataset::FlushSlot(PushSlot* s) {
  if (s->seg_rows == 0) return;
  LGB_CUDA(cudaSetDevice(device));
  const int k = s->cur;
  const size_t bytes = static_cast<size_t>(s->seg_rows) * num_total_features * sizeof(double);
  LGB_CUDA(cudaMemcpyAsync(s->dstage[k], s->pinned[k], bytes, cudaMemcpyHostToDevice, s->stream));
  const int F = num_features();
  if (F > 0) {
    LaunchBinRows(s->dstage[k], static_cast<int>(s->seg_rows), s->seg_start, s->stream);
    LGB_CUDA(cudaGetLastError());
  }
  LGB_CUDA(cudaEventRecord(s->ev[k], s->stream));
  s->cur ^= 1;
  LGB_CUDA(cudaEventSynchronize(s->ev[s->cur]));   // the buffer we are about to refill must have been consumed
  s->seg_rows = 0;
}

// Measurement hook (bench.py's K2 line): `nrow` rows of raw values are copied to the device once (untimed) and the production
// binning kernel is timed, with CUDA events, re-binning the dataset's first `nrow` rows `repeats` times from that resident copy.
// The rows must be the dataset's own first rows, so the bins that are rewritten do not change.
float Dataset::TimeBinning(const double* rows, int nrow, int repeats) {
  LGB_CHECK(finished, "TimeBinning: the dataset must be finished");
  LGB_CHECK(nrow > 0 && nrow <= num_data && repeats > 0, "TimeBinning: bad arguments");
  LGB_CUDA(cudaSetDevice(device));
  const int F = num_features();
  LGB_CHECK(F > 0, "TimeBinning: no used feature");
  DevBuf<double> d;
  d.Alloc(static_cast<size_t>(nrow) * num_total_features);
  LGB_CUDA(cudaMemcpy(d.p, rows, sizeof(double) * nrow * num_total_features, cudaMemcpyHostToDevice));
  cudaEvent_t e0, e1;
  LGB_CUDA(cudaEventCreate(&e0)); LGB_CUDA(cudaEventCreate(&e1));
  LaunchBinRows(d.p, nrow, 0, 0);
  LGB_CUDA(cudaEventRecord(e0, 0));
  for (int r = 0; r < repeats; ++r) LaunchBinRows(d.p, nrow, 0, 0);
  LGB_CUDA(cudaEventRecord(e1, 0));
  LGB_CUDA(cudaEventSynchronize(e1));
  LGB_CUDA(cudaGetLastError());
  float ms = 0;
  LGB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  return ms / repeats;
}

void Dataset::PushRows(const void* data, int data_type, int32_t nrow, int32_t ncol, int64_t start_row, const float* lab,
                       const float* w, const double* init, const int32_t* q, int tid) {
  LGB_CHECK(data_type == 1 || data_type == 0, "PushRows: data_type must be C_API_DTYPE_FLOAT32 or FLOAT64");
  LGB_CHECK(ncol == num_total_features, "PushRows: got %d columns, the dataset has %d features", ncol, num_total_features);
  LGB_CHECK(nrow >= 0 && start_row >= 0 && start_row + nrow <= num_data, "PushRows: rows [%lld, %lld) are outside the dataset (%lld rows)",
            (long long)start_row, (long long)(start_row + nrow), (long long)num_data);
  LGB_CHECK(!finished, "PushRows: the dataset is already finished");
  if (tid < 0) tid = 0;
  if (static_cast<int>(slots_.size()) <= tid) EnsureSlots(tid + 1);
  PushSlot* s = &slots_[tid];
  int64_t done = 0;
  while (done < nrow) {
    if (s->seg_rows > 0 && (start_row + done != s->seg_start + s->seg_rows || static_cast<size_t>(s->seg_rows) >= s->cap_rows)) FlushSlot(s);
    if (s->seg_rows == 0) s->seg_start = start_row + done;
    const int64_t take = std::min<int64_t>(nrow - done, static_cast<int64_t>(s->cap_rows) - s->seg_rows);
    double* dst = s->pinned[s->cur] + static_cast<size_t>(s->seg_rows) * ncol;
    if (data_type == 1) {
      memcpy(dst, static_cast<const double*>(data) + static_cast<size_t>(done) * ncol, sizeof(double) * take * ncol);
    } else {
      const float* src = static_cast<const float*>(data) + static_cast<size_t>(done) * ncol;
      for (size_t i = 0; i < static_cast<size_t>(take) * ncol; ++i) dst[i] = src[i];
    }
    s->seg_rows += take;
    done += take;
  }
  PushMetadata(nrow, start_row, lab, w, init, q);
}

// Sparse micro-batches (StreamingPartitionTask.scala:239-277): the CSR rows are expanded into the same pinned dense staging rows
// the dense path fills (zero fill + scatter, a formatting step at memory speed), then take the dense path: H2D and the K2 binning
// kernel.  Absent entries are 0.0, exactly what the reference's SparseVector means.
void Dataset::PushRowsByCSR(const void* indptr, int indptr_type, const int32_t* indices, const void* data, int data_type,
                            int64_t nindptr, int64_t nelem, int64_t start_row, const float* lab, const float* w, const double* init,
                            const int32_t* q, int tid) {
  LGB_CHECK(indptr_type == 2 || indptr_type == 3, "PushRowsByCSR: indptr_type must be C_API_DTYPE_INT32 or INT64");
  LGB_CHECK(data_type == 1 || data_type == 0, "PushRowsByCSR: data_type must be C_API_DTYPE_FLOAT32 or FLOAT64");
  const int64_t nrow = nindptr - 1;
  const int ncol = num_total_features;
  LGB_CHECK(nrow >= 0 && start_row >= 0 && start_row + nrow <= num_data, "PushRowsByCSR: rows [%lld, %lld) are outside the dataset (%lld rows)",
            (long long)start_row, (long long)(start_row + nrow), (long long)num_data);
  LGB_CHECK(!finished, "PushRowsByCSR: the dataset is already finished");
  auto ptr_at = [&](int64_t i) -> int64_t {
    return indptr_type == 3 ? static_cast<const int64_t*>(indptr)[i] : static_cast<int64_t>(static_cast<const int32_t*>(indptr)[i]);
  };
  LGB_CHECK(nrow == 0 || (ptr_at(nrow) - ptr_at(0) <= nelem && ptr_at(0) >= 0), "PushRowsByCSR: indptr does not match nelem");
  if (tid < 0) tid = 0;
  if (static_cast<int>(slots_.size()) <= tid) EnsureSlots(tid + 1);
  PushSlot* s = &slots_[tid];
  int64_t done = 0;
  while (done < nrow) {
    if (s->seg_rows > 0 && (start_row + done != s->seg_start + s->seg_rows || static_cast<size_t>(s->seg_rows) >= s->cap_rows)) FlushSlot(s);
    if (s->seg_rows == 0) s->seg_start = start_row + done;
    const int64_t take = std::min<int64_t>(nrow - done, static_cast<int64_t>(s->cap_rows) - s->seg_rows);
    double* dst = s->pinned[s->cur] + static_cast<size_t>(s->seg_rows) * ncol;
    memset(dst, 0, sizeof(double) * take * ncol);
    for (int64_t r = 0; r < take; ++r) {
      double* row = dst + static_cast<size_t>(r) * ncol;
      for (int64_t j = ptr_at(done + r); j < ptr_at(done + r + 1); ++j) {
        const int c = indices[j];
        LGB_CHECK(c >= 0 && c < ncol, "PushRowsByCSR: column index %d is outside [0, %d)", c, ncol);
        row[c] = data_type == 1 ? static_cast<const double*>(data)[j] : static_cast<double>(static_cast<const float*>(data)[j]);
      }
    }
    s->seg_rows += take;
    done += take;
  }
  PushMetadata(static_cast<int32_t>(nrow), start_row, lab, w, init, q);
}

void Dataset::PushMetadata(int32_t nrow, int64_t start_row, const float* lab, const float* w, const double* init, const int32_t* q) {
  if (lab) std::copy(lab, lab + nrow, label.begin() + start_row);
  if (w && has_weight) std::copy(w, w + nrow, weight.begin() + start_row);
  if (init && has_init_score) {
    // a micro-batch carries its init scores class-major, [nclasses][nrow] (StreamingPartitionTask.scala:213-219)
    for (int k = 0; k < num_init_score_classes; ++k)
      std::copy(init + static_cast<size_t>(k) * nrow, init + static_cast<size_t>(k + 1) * nrow, init_score.begin() + static_cast<size_t>(k) * num_data + start_row);
  }
  if (q && has_query) std::copy(q, q + nrow, query_id.begin() + start_row);
  bool all_in = false;
  {
    std::lock_guard<std::mutex> g(mu_);
    pushed_rows += nrow;
    all_in = pushed_rows >= num_data;
  }
  if (all_in && !wait_for_manual_finish) MarkFinished();
}

void Dataset::MarkFinished() {
  if (finished) return;
  LGB_CUDA(cudaSetDevice(device));
  for (auto& s : slots_) FlushSlot(&s);
  for (auto& s : slots_) LGB_CUDA(cudaStreamSynchronize(s.stream));
  FreeSlots();
  d_label.Alloc(std::max<int64_t>(num_data, 1));
  LGB_CUDA(cudaMemcpy(d_label.p, label.data(), sizeof(float) * num_data, cudaMemcpyHostToDevice));
  if (has_weight) {
    d_weight.Alloc(std::max<int64_t>(num_data, 1));
    LGB_CUDA(cudaMemcpy(d_weight.p, weight.data(), sizeof(float) * num_data, cudaMemcpyHostToDevice));
  }
  if (has_init_score) {
    d_init_score.Alloc(std::max<size_t>(init_score.size(), 1));
    LGB_CUDA(cudaMemcpy(d_init_score.p, init_score.data(), sizeof(double) * init_score.size(), cudaMemcpyHostToDevice));
  }
  if (has_query && query_boundaries.empty()) {
    // runs of equal consecutive group ids are queries (the streaming path pushes one id per row)
    query_boundaries.push_back(0);
    for (int64_t i = 1; i < num_data; ++i)
      if (query_id[i] != query_id[i - 1]) query_boundaries.push_back(static_cast<int32_t>(i));
    query_boundaries.push_back(static_cast<int32_t>(num_data));
  }
  if (!query_boundaries.empty()) {
    d_query_boundaries.Alloc(query_boundaries.size());
    LGB_CUDA(cudaMemcpy(d_query_boundaries.p, query_boundaries.data(), sizeof(int32_t) * query_boundaries.size(), cudaMemcpyHostToDevice));
  }
  finished = true;
}

void Dataset::CopyBinColumn(int logical_feature, uint16_t* out) const {
  LGB_CHECK(logical_feature >= 0 && logical_feature < num_logical_features(), "feature %d out of range", logical_feature);
  LGB_CUDA(cudaSetDevice(device));
  const int first = logical_first_col[logical_feature], cols = logical_num_cols[logical_feature];
  std::vector<uint8_t> col(num_data);
  std::vector<int> seen(num_data, 0);
  for (int c = 0; c < cols; ++c) {
    LGB_CUDA(cudaMemcpy(col.data(), d_cm.p + static_cast<size_t>(first + c) * n_pad, num_data, cudaMemcpyDeviceToHost));
    for (int64_t i = 0; i < num_data; ++i) {
      if (cols == 1) { out[i] = col[i]; continue; }
      if (col[i] >= 1 && col[i] <= kWideLocalBins) { out[i] = static_cast<uint16_t>(c * kWideLocalBins + col[i] - 1); ++seen[i]; }
    }
  }
  if (cols > 1)
    for (int64_t i = 0; i < num_data; ++i) LGB_CHECK(seen[i] == 1, "row %lld of a wide feature is coded in %d columns", (long long)i, seen[i]);
}

void Dataset::CopyRowMajor(int64_t row, int64_t nrows, uint8_t* out) const {
  LGB_CHECK(row >= 0 && row + nrows <= num_data, "row range out of bounds");
  LGB_CUDA(cudaSetDevice(device));
  LGB_CUDA(cudaMemcpy(out, d_rm.p + static_cast<size_t>(row) * layout.rm_stride, static_cast<size_t>(nrows) * layout.rm_stride, cudaMemcpyDeviceToHost));
}

}  // namespace lgb
