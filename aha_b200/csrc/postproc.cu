// Highlight post-processing on the score stream: find_ticks of test/live_infer_for_video.py:195-228 and
// test/live_video/infer_live_video.py:70-91, i.e. scipy.signal.savgol_filter(x, 15, 3) -> thresh = mean + 0.5*std ->
// scipy.signal.find_peaks(height=thresh, prominence, distance), as ONE kernel: one 256-thread block per stream, float64
// throughout (the reference works on a Python float list -> np.float64).  The smoothed signal, the candidate peaks and
// their state stay in shared memory (two blocks per SM at 10 000 samples); mean and std are added up in numpy's pairwise
// order, so the height threshold is bit-identical to the reference's.
//
// Algorithmic bytes per stream: read T*8 B of scores, write the peak list (+ T*8 B smoothed when asked for).
// The find_peaks stages are exact restatements of scipy/signal/_peak_finding_utils.pyx:
//   local maxima with plateau midpoints -> height -> distance (greedy by descending height, here as an equivalent
//   parallel fixed point: an undecided peak with no higher undecided peak within `distance` is kept, then everything
//   within `distance` of a kept peak is dropped, until nothing is undecided) -> prominence over the full signal.
#include "ops.h"
#include "savgol_tables.h"
#include <cstdio>
#include <cstdlib>

namespace {

__constant__ double cF[15];
__constant__ double cHead[7][15];
__constant__ double cTail[7][15];

constexpr int kPPThreads = 256;

// exclusive block scan of one int per thread; returns the exclusive prefix, *total gets the block sum
__device__ __forceinline__ int block_exscan(int v, int* sh, int* total) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  int inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { const int n = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += n; }
  __syncthreads();
  if (lane == 31) sh[w] = inc;
  __syncthreads();
  int base = 0, tot = 0;
  for (int i = 0; i < kPPThreads / 32; ++i) { if (i < w) base += sh[i]; tot += sh[i]; }
  *total = tot;
  return base + inc - v;
}

// ---- numpy's pairwise summation (numpy/_core/src/umath/loops_utils.h.src, DOUBLE_pairwise_sum), which is what
// smoothed.mean() and smoothed.std() of the reference add up with: blocks of at most 128 elements are summed with eight
// interleaved accumulators, larger ranges are split at n/2 rounded down to a multiple of 8 and the halves added.
// Reproducing the association order makes mean, std and therefore the height threshold BIT-IDENTICAL to numpy's, so a peak
// whose smoothed height sits exactly on the threshold is kept or dropped exactly as scipy does.
struct PwLeaf { int start, len; };

// leaf table of the recursion over n elements, in left-to-right order (one thread); returns the leaf count
__device__ int pw_build_leaves(int n, PwLeaf* leaves, int cap) {
  int stack_start[48], stack_len[48];
  int sp = 0, nl = 0;
  stack_start[0] = 0; stack_len[0] = n; sp = 1;
  while (sp > 0) {
    --sp;
    const int st = stack_start[sp], ln = stack_len[sp];
    if (ln <= 128) { if (nl < cap) leaves[nl] = PwLeaf{st, ln}; ++nl; continue; }
    int n2 = ln / 2; n2 -= n2 % 8;
    // right half first onto the stack so that the left half is expanded first (leaves come out in order)
    stack_start[sp] = st + n2; stack_len[sp] = ln - n2; ++sp;
    stack_start[sp] = st; stack_len[sp] = n2; ++sp;
  }
  return nl;
}

template <bool kCentered>
__device__ __forceinline__ double pw_elem(const double* y, int i, double mean) {
  if (!kCentered) return y[i];
  const double d = __dsub_rn(y[i], mean);
  return __dmul_rn(d, d);
}

// The same leaf sum by EIGHT lanes (lane j of the group owns accumulator r[j] of numpy's unrolled loop, i.e. the elements
// j, j + 8, ...): consecutive lanes read consecutive elements (one thread per leaf walked 1 KB apart: a 32-way bank
// conflict) and the dependent chain is 16 additions instead of 128.  Result in every lane of the group.
template <bool kCentered>
__device__ __forceinline__ double pw_leaf_sum8(const double* y, PwLeaf lf, double mean, int sub) {
  const int o = lf.start, n = lf.len;
  double res;
  if (n < 8) {
    res = 0.0;
    for (int i = 0; i < n; ++i) res = __dadd_rn(res, pw_elem<kCentered>(y, o + i, mean));
    return res;
  }
  const int n8 = n - (n % 8);
  double r = pw_elem<kCentered>(y, o + sub, mean);
  for (int i = 8 + sub; i < n8; i += 8) r = __dadd_rn(r, pw_elem<kCentered>(y, o + i, mean));
  // ((r0 + r1) + (r2 + r3)) + ((r4 + r5) + (r6 + r7)): xor-butterfly over the 8 lanes adds exactly these pairs
  r = __dadd_rn(r, __shfl_xor_sync(0xffffffffu, r, 1));
  r = __dadd_rn(r, __shfl_xor_sync(0xffffffffu, r, 2));
  r = __dadd_rn(r, __shfl_xor_sync(0xffffffffu, r, 4));
  res = r;
  for (int i = n8; i < n; ++i) res = __dadd_rn(res, pw_elem<kCentered>(y, o + i, mean));
  return res;
}

// adds the leaf sums in the recursion's order (one thread): same splitting rule, leaves consumed left to right.  Iterative
// (explicit node and value stacks): the recursive form cost ~250 cycles per call in local-memory frames.
__device__ double pw_combine(const double* leaf_sums, int n) {
  int node_n[64];
  unsigned char node_post[64];
  double val[64];
  int sp = 0, vp = 0, li = 0;
  node_n[0] = n; node_post[0] = 0; sp = 1;
  while (sp > 0) {
    --sp;
    const int ln = node_n[sp];
    if (node_post[sp]) { const double b = val[--vp]; const double a = val[--vp]; val[vp++] = __dadd_rn(a, b); continue; }
    if (ln <= 128) { val[vp++] = leaf_sums[li++]; continue; }
    int n2 = ln / 2; n2 -= n2 % 8;
    node_n[sp] = ln; node_post[sp] = 1; ++sp;            // add the two halves once both are on the value stack
    node_n[sp] = ln - n2; node_post[sp] = 0; ++sp;       // right half (popped second)
    node_n[sp] = n2; node_post[sp] = 0; ++sp;            // left half (popped first)
  }
  return val[0];
}

// IEEE doubles ordered as unsigned 64-bit integers (sign bit flipped for non-negative values, all bits for negative ones):
// every comparison of the peak logic runs on these keys - the FP64 pipe of the B200's CUDA cores issues a warp instruction
// only every ~16 cycles per scheduler (measured: the all-double version of this kernel was bound by DSETP / DMNMX / DADD
// issue, 0.4 ms per 10 000-sample stream), 64-bit integer compares cost two ALU instructions.
__device__ __forceinline__ unsigned long long dkey(double v) {
  const unsigned long long b = (unsigned long long)__double_as_longlong(v);
  return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}
__device__ __forceinline__ double dkey_inv(unsigned long long k) {
  const unsigned long long b = (k >> 63) ? (k & 0x7fffffffffffffffull) : ~k;
  return __longlong_as_double((long long)b);
}
__device__ __forceinline__ unsigned long long warp_min_u64(unsigned long long v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const unsigned long long w = __shfl_xor_sync(0xffffffffu, v, o);
    v = w < v ? w : v;
  }
  return v;
}

// kSmem: the smoothed signal, the candidate list and its state live in shared memory (T up to ~21 000 samples); otherwise
// in the per-stream global workspaces.  Algorithmic bytes per stream: T * 8 read, the peak list written.
template <bool kSmem>
struct PPIdx { typedef int32_t type; };
template <>
struct PPIdx<true> { typedef uint16_t type; };

template <bool kSmem>
__global__ void __launch_bounds__(kPPThreads) find_ticks_kernel(const double* __restrict__ scores, int T, double prominence,
                                                               int distance, const double* __restrict__ thresh_in,
                                                               int32_t* __restrict__ peaks_out, int max_peaks,
                                                               int32_t* __restrict__ n_peaks, double* __restrict__ smoothed,
                                                               double* __restrict__ thresh_out, double* __restrict__ y_ws,
                                                               int32_t* __restrict__ cand_ws, int8_t* __restrict__ state_ws,
                                                               int cand_cap, int leaf_cap, int dbg) {
  extern __shared__ __align__(16) uint8_t pp_smem[];
  __shared__ int sh_i[kPPThreads / 32];
  __shared__ int sh_nleaves;
  __shared__ double sh_val;
  const int s = blockIdx.x;
  const double* x = scores + (size_t)s * T;
  // shared-memory carve-up: leaf table + leaf sums always; y / cand / state when they fit
  PwLeaf* leaves = reinterpret_cast<PwLeaf*>(pp_smem);
  double* leaf_sums = reinterpret_cast<double*>(pp_smem + (size_t)leaf_cap * sizeof(PwLeaf));
  uint8_t* after = reinterpret_cast<uint8_t*>(leaf_sums + leaf_cap);
  double* y = kSmem ? reinterpret_cast<double*>(after) : y_ws + (size_t)s * T;
  // candidate positions and their height order: 16-bit in shared memory (T < 32768 there), 32-bit in the global fallback
  typedef typename PPIdx<kSmem>::type idx_t;
  idx_t* cand = kSmem ? reinterpret_cast<idx_t*>(after + (size_t)T * 8)
                      : reinterpret_cast<idx_t*>(cand_ws + (size_t)s * 2 * cand_cap);
  idx_t* ord = cand + cand_cap;
  int8_t* state = kSmem ? reinterpret_cast<int8_t*>(after + (size_t)T * 8 + (size_t)cand_cap * 2 * sizeof(idx_t))
                        : state_ws + (size_t)s * cand_cap;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  long long t_[8];      // bring-up (AHA_PP_DBG=1): stage stamps of block 0
  t_[0] = clock64();

  // ---- A. Savitzky-Golay(15, 3), mode='interp' ----
  for (int i = tid; i < T; i += kPPThreads) {
    double v;
    if (i >= 7 && i < T - 7) {
      // scipy.ndimage.correlate1d, symmetric-filter branch: tmp = x[i]*f[7]; tmp += (x[i+l] + x[i-l]) * f[l+7], l=-7..-1
      v = __dmul_rn(x[i], cF[7]);
#pragma unroll
      for (int l = -7; l < 0; ++l) v = __dadd_rn(v, __dmul_rn(__dadd_rn(x[i + l], x[i - l]), cF[l + 7]));
    } else if (i < 7) {
      v = 0.0;
      for (int k = 0; k < 15; ++k) v = __dadd_rn(v, __dmul_rn(cHead[i][k], x[k]));
    } else {
      v = 0.0;
      const int r = i - (T - 7);
      for (int k = 0; k < 15; ++k) v = __dadd_rn(v, __dmul_rn(cTail[r][k], x[T - 15 + k]));
    }
    y[i] = v;
    if (kSmem && smoothed) smoothed[(size_t)s * T + i] = v;
  }
  if (tid == 0) sh_nleaves = pw_build_leaves(T, leaves, leaf_cap);
  __syncthreads();

  t_[1] = clock64();
  // ---- B. threshold = mean + 0.5 * std (ddof = 0) of the smoothed signal, in numpy's summation order ----
  double th;
  const bool auto_thresh = (thresh_in == nullptr) || isnan(thresh_in[s]);
  if (auto_thresh) {
    const int nl = min(sh_nleaves, leaf_cap);
    // (all 32 lanes of a warp run the same trip count: the leaf index is clamped, surplus groups do not store)
    for (int l0 = 0; l0 < nl; l0 += kPPThreads / 8) {
      const int l = l0 + (tid >> 3);
      const double v = pw_leaf_sum8<false>(y, leaves[min(l, nl - 1)], 0.0, tid & 7);
      if (l < nl && (tid & 7) == 0) leaf_sums[l] = v;
    }
    __syncthreads();
    if (tid == 0) sh_val = __ddiv_rn(pw_combine(leaf_sums, T), (double)T);
    __syncthreads();
    const double mean = sh_val;
    for (int l0 = 0; l0 < nl; l0 += kPPThreads / 8) {
      const int l = l0 + (tid >> 3);
      const double v = pw_leaf_sum8<true>(y, leaves[min(l, nl - 1)], mean, tid & 7);
      if (l < nl && (tid & 7) == 0) leaf_sums[l] = v;
    }
    __syncthreads();
    if (tid == 0) {
      const double var = __ddiv_rn(pw_combine(leaf_sums, T), (double)T);
      sh_val = __dadd_rn(mean, __dmul_rn(0.5, sqrt(var)));
    }
    __syncthreads();
    th = sh_val;
  } else {
    th = thresh_in[s];
  }
  if (tid == 0 && thresh_out) thresh_out[s] = th;

  __syncthreads();
  // from here on the signal is only compared: keep it as orderable integer keys (in place; -0.0 == +0.0 is the one
  // equality the keys would miss, so zeros are normalised)
  unsigned long long* yk = reinterpret_cast<unsigned long long*>(y);
  for (int i = tid; i < T; i += kPPThreads) { const double v = y[i]; yk[i] = dkey(v == 0.0 ? 0.0 : v); }
  const unsigned long long thk = dkey(th == 0.0 ? 0.0 : th);
  __syncthreads();
  t_[2] = clock64();
  // ---- C. strict local maxima with plateau midpoints, filtered by height, compacted in position order ----
  const int seg = ((T + kPPThreads - 1) / kPPThreads) | 1;   // odd: the lanes' 8-byte reads then fall into distinct banks
  const int i0 = max(1, tid * seg), i1 = min(T - 1, (tid + 1) * seg);
  int cnt = 0;
  for (int pass = 0; pass < 2; ++pass) {
    int wr = 0, total = 0;
    if (pass == 1) wr = block_exscan(cnt, sh_i, &total);
    for (int i = i0; i < i1; ++i) {
      if (yk[i - 1] < yk[i]) {
        int a = i + 1;
        while (a < T - 1 && yk[a] == yk[i]) ++a;
        if (yk[a] < yk[i]) {
          const int pk = (i + a - 1) / 2;
          if (yk[pk] >= thk) {
            if (pass == 0) ++cnt;
            else { if (wr < cand_cap) { cand[wr] = (idx_t)pk; state[wr] = 1; } ++wr; }
          }
        }
      }
    }
    if (pass == 1) cnt = total;
  }
  __syncthreads();
  const int m = min(cnt, cand_cap);

  t_[3] = clock64();
  // ---- D. distance (scipy _select_by_peak_distance): peaks are visited from the highest down, a visited peak that is
  //         still kept removes every peak closer than `distance`.  The visiting order comes from a rank count (m^2 / 256
  //         comparisons per thread; equal heights: the larger index first), the greedy pass itself is sequential - a few
  //         hundred steps on shared memory.  (The first version iterated a parallel fixed point: one round per link of the
  //         longest chain of ever higher neighbours, a hundred rounds of four barriers on a noisy slope.) ----
  if (distance > 1) {
    for (int k = tid; k < m; k += kPPThreads) {
      const unsigned long long hk = yk[cand[k]];
      int rank = 0;
      for (int j = 0; j < m; ++j) {
        const unsigned long long hj = yk[cand[j]];
        rank += (hj > hk) || (hj == hk && j > k);
      }
      ord[rank] = (idx_t)k;
    }
    __syncthreads();
    if (warp == 0) {
      // the greedy pass, one warp: the lanes clear the neighbours of a kept peak 32 at a time
      for (int r = 0; r < m; ++r) {
        const int j = ord[r];
        if (state[j] != 1) continue;                               // (warp-uniform: shared memory, ordered by __syncwarp)
        const int pj = cand[j];
        for (int k0 = j - 1; k0 >= 0; k0 -= 32) {
          const int k = k0 - lane;
          const bool in = k >= 0 && pj - (int)cand[k] < distance;
          if (in) state[k] = 2;
          if (!__all_sync(0xffffffffu, in)) break;
        }
        for (int k0 = j + 1; k0 < m; k0 += 32) {
          const int k = k0 + lane;
          const bool in = k < m && (int)cand[k] - pj < distance;
          if (in) state[k] = 2;
          if (!__all_sync(0xffffffffu, in)) break;
        }
        __syncwarp();
      }
    }
    __syncthreads();
  }

  t_[4] = clock64();
  // ---- E. prominence over the full signal (wlen=None): one WARP per kept peak walks outwards 32 samples at a time
  //         (the walk of a high peak crosses most of the signal: one thread per peak took a dependent load per sample) ----
  for (int k = warp; k < m; k += kPPThreads / 32) {
    if (state[k] != 1) continue;                                  // (warp-uniform)
    const int pk = cand[k];
    const unsigned long long h = yk[pk];
    unsigned long long lmin = h, rmin = h;
    for (int base = pk; base >= 0; base -= 32) {                  // left: i = pk, pk-1, ... while y[i] <= h
      const int i = base - lane;
      const unsigned long long v = (i >= 0) ? yk[i] : h;
      const uint32_t higher = __ballot_sync(0xffffffffu, i >= 0 && v > h);
      const int stop = higher ? __ffs(higher) - 1 : 32;           // first lane (nearest sample) that is strictly higher
      const unsigned long long mn = warp_min_u64(lane < stop ? v : h);
      lmin = mn < lmin ? mn : lmin;
      if (higher) break;
    }
    for (int base = pk; base < T; base += 32) {                   // right
      const int i = base + lane;
      const unsigned long long v = (i < T) ? yk[i] : h;
      const uint32_t higher = __ballot_sync(0xffffffffu, i < T && v > h);
      const int stop = higher ? __ffs(higher) - 1 : 32;
      const unsigned long long mn = warp_min_u64(lane < stop ? v : h);
      rmin = mn < rmin ? mn : rmin;
      if (higher) break;
    }
    // scipy: prominence = x[peak] - max(left_min, right_min), kept if >= the bound (one double subtraction per peak)
    if (lane == 0 && !(__dsub_rn(dkey_inv(h), dkey_inv(lmin > rmin ? lmin : rmin)) >= prominence)) state[k] = 2;
  }
  __syncthreads();
  t_[5] = clock64();
  const int cseg = (m + kPPThreads - 1) / kPPThreads;
  const int k0 = min(m, tid * cseg), k1 = min(m, (tid + 1) * cseg);
  int mine = 0;
  for (int k = k0; k < k1; ++k) mine += (state[k] == 1);
  int total = 0;
  int wr = block_exscan(mine, sh_i, &total);
  for (int k = k0; k < k1; ++k)
    if (state[k] == 1) { if (wr < max_peaks) peaks_out[(size_t)s * max_peaks + wr] = (int32_t)cand[k]; ++wr; }
  if (tid == 0) n_peaks[s] = (cnt > cand_cap) ? 0x7fffffff : total;
  if (!kSmem && smoothed) {      // (the global path kept the signal in the caller's `smoothed` buffer: back to doubles)
    __syncthreads();
    for (int i = tid; i < T; i += kPPThreads) y[i] = dkey_inv(yk[i]);
  }
  if (dbg && tid == 0 && s == 0)
    printf("find_ticks block 0: m=%d kept=%d cycles A %lld B %lld C %lld D %lld E %lld F %lld\n", m, total, t_[1] - t_[0], t_[2] - t_[1],
           t_[3] - t_[2], t_[4] - t_[3], t_[5] - t_[4], clock64() - t_[5]);
}

bool g_tables_loaded[16] = {false};

}  // namespace

int op_find_ticks(const double* scores, int n_streams, int T, double prominence, int distance, const double* thresh,
                  int32_t* peaks, int max_peaks, int32_t* n_peaks, double* smoothed, double* thresh_out,
                  cudaStream_t st) {
  AHA_REQUIRE(T >= 15, "If mode is 'interp', window_length must be less than or equal to the size of x.");
  AHA_REQUIRE(distance >= 1, "`distance` must be greater or equal to 1");
  AHA_REQUIRE(n_streams > 0 && max_peaks > 0 && scores && peaks && n_peaks, "find_ticks: bad arguments");
  int dev = 0;
  AHA_CUDA(cudaGetDevice(&dev));
  if (dev < 16 && !g_tables_loaded[dev]) {
    AHA_CUDA(cudaMemcpyToSymbol(cF, kSavgolF, sizeof(kSavgolF)));
    AHA_CUDA(cudaMemcpyToSymbol(cHead, kSavgolHead, sizeof(kSavgolHead)));
    AHA_CUDA(cudaMemcpyToSymbol(cTail, kSavgolTail, sizeof(kSavgolTail)));
    g_tables_loaded[dev] = true;
  }
  static const int dbg = getenv("AHA_PP_DBG") ? atoi(getenv("AHA_PP_DBG")) : 0;
  const int cand_cap = T / 2 + 1;  // strict local maxima are at least 2 samples apart
  const int leaf_cap = T / 48 + 8; // leaves of numpy's pairwise recursion hold 57 .. 128 elements
  const size_t base_smem = (size_t)leaf_cap * (sizeof(int) * 2 + sizeof(double));
  const size_t full_smem = base_smem + (size_t)T * 8 + (size_t)cand_cap * 5 + 16;      // y, cand + ord (16-bit), state
  const bool in_smem = full_smem <= 225 * 1024 && T < 32768;
  double* y_ws = nullptr;
  int32_t* cand_ws = nullptr;
  int8_t* state_ws = nullptr;
  if (!in_smem) {
    if (!smoothed) AHA_CUDA(cudaMallocAsync(&y_ws, (size_t)n_streams * T * sizeof(double), st));
    AHA_CUDA(cudaMallocAsync(&cand_ws, (size_t)n_streams * 2 * cand_cap * sizeof(int32_t), st));     // positions + height order
    AHA_CUDA(cudaMallocAsync(&state_ws, (size_t)n_streams * cand_cap, st));
    AHA_REQUIRE(base_smem <= 200 * 1024, "find_ticks: signal too long");
    static bool attr_g[16] = {false};
    if (dev < 16 && !attr_g[dev]) {
      AHA_CUDA(cudaFuncSetAttribute(find_ticks_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
      attr_g[dev] = true;
    }
    find_ticks_kernel<false><<<n_streams, kPPThreads, base_smem, st>>>(scores, T, prominence, distance, thresh, peaks, max_peaks,
                                                                      n_peaks, smoothed, thresh_out, smoothed ? smoothed : y_ws,
                                                                      cand_ws, state_ws, cand_cap, leaf_cap, dbg);
  } else {
    static bool attr_s[16] = {false};
    if (dev < 16 && !attr_s[dev]) {
      AHA_CUDA(cudaFuncSetAttribute(find_ticks_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 225 * 1024));
      attr_s[dev] = true;
    }
    find_ticks_kernel<true><<<n_streams, kPPThreads, full_smem, st>>>(scores, T, prominence, distance, thresh, peaks, max_peaks,
                                                                     n_peaks, smoothed, thresh_out, nullptr, nullptr, nullptr,
                                                                     cand_cap, leaf_cap, dbg);
  }
  ++g_aha_launch_count;
  cudaError_t e = cudaGetLastError();
  if (y_ws) cudaFreeAsync(y_ws, st);
  if (cand_ws) cudaFreeAsync(cand_ws, st);
  if (state_ws) cudaFreeAsync(state_ws, st);
  if (e != cudaSuccess) return aha_cuda_fail(e, "find_ticks_kernel", __FILE__, __LINE__);
  return AHA_OK;
}
