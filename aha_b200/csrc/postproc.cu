// Highlight post-processing on the score stream: find_ticks of test/live_infer_for_video.py:195-228 and
// test/live_video/infer_live_video.py:70-91, i.e. scipy.signal.savgol_filter(x, 15, 3) -> thresh = mean + 0.5*std ->
// scipy.signal.find_peaks(height=thresh, prominence, distance), as ONE kernel: one 256-thread block per stream, float64
// throughout (the reference works on a Python float list -> np.float64).
//
// Algorithmic bytes per stream: read T*8 B of scores, write the peak list (+ T*8 B smoothed when asked for).
// The find_peaks stages are exact restatements of scipy/signal/_peak_finding_utils.pyx:
//   local maxima with plateau midpoints -> height -> distance (greedy by descending height, here as an equivalent
//   parallel fixed point: an undecided peak with no higher undecided peak within `distance` is kept, then everything
//   within `distance` of a kept peak is dropped, until nothing is undecided) -> prominence over the full signal.
#include "ops.h"
#include "savgol_tables.h"

namespace {

__constant__ double cF[15];
__constant__ double cHead[7][15];
__constant__ double cTail[7][15];

constexpr int kPPThreads = 256;

__device__ __forceinline__ double block_sum_d(double v, double* red) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if (lane == 0) red[w] = v;
  __syncthreads();
  double r = 0.0;
  for (int i = 0; i < kPPThreads / 32; ++i) r += red[i];
  return r;
}

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

__global__ void __launch_bounds__(kPPThreads) find_ticks_kernel(const double* __restrict__ scores, int T, double prominence,
                                                               int distance, const double* __restrict__ thresh_in,
                                                               int32_t* __restrict__ peaks_out, int max_peaks,
                                                               int32_t* __restrict__ n_peaks, double* __restrict__ smoothed,
                                                               double* __restrict__ thresh_out, int32_t* __restrict__ cand_ws,
                                                               int8_t* __restrict__ state_ws, int cand_cap) {
  __shared__ double red[kPPThreads / 32];
  __shared__ int sh_i[kPPThreads / 32];
  __shared__ int sh_flag;
  const int s = blockIdx.x;
  const double* x = scores + (size_t)s * T;
  double* y = smoothed + (size_t)s * T;
  int32_t* cand = cand_ws + (size_t)s * cand_cap;
  int8_t* state = state_ws + (size_t)s * cand_cap;
  const int tid = threadIdx.x;

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
  }
  __syncthreads();

  // ---- B. threshold = mean + 0.5 * std (ddof = 0) of the smoothed signal ----
  double th;
  const bool auto_thresh = (thresh_in == nullptr) || isnan(thresh_in[s]);
  {
    double acc = 0.0;
    for (int i = tid; i < T; i += kPPThreads) acc += y[i];
    const double mean = block_sum_d(acc, red) / (double)T;
    acc = 0.0;
    for (int i = tid; i < T; i += kPPThreads) { const double dl = y[i] - mean; acc += dl * dl; }
    const double var = block_sum_d(acc, red) / (double)T;
    th = auto_thresh ? mean + 0.5 * sqrt(var) : thresh_in[s];
  }
  if (tid == 0 && thresh_out) thresh_out[s] = th;

  // ---- C. strict local maxima with plateau midpoints, filtered by height, compacted in position order ----
  const int seg = (T + kPPThreads - 1) / kPPThreads;
  const int i0 = max(1, tid * seg), i1 = min(T - 1, (tid + 1) * seg);
  int cnt = 0;
  for (int pass = 0; pass < 2; ++pass) {
    int wr = 0, total = 0;
    if (pass == 1) wr = block_exscan(cnt, sh_i, &total);
    for (int i = i0; i < i1; ++i) {
      if (y[i - 1] < y[i]) {
        int a = i + 1;
        while (a < T - 1 && y[a] == y[i]) ++a;
        if (y[a] < y[i]) {
          const int pk = (i + a - 1) / 2;
          if (y[pk] >= th) {
            if (pass == 0) ++cnt;
            else { if (wr < cand_cap) { cand[wr] = pk; state[wr] = 0; } ++wr; }
          }
        }
      }
    }
    if (pass == 1) cnt = total;
  }
  __syncthreads();
  const int m = min(cnt, cand_cap);

  // ---- D. distance: parallel fixed point of the greedy highest-first suppression ----
  if (distance > 1) {
    for (int round = 0; round < m + 1; ++round) {
      if (tid == 0) sh_flag = 0;
      __syncthreads();
      // phase 1: undecided peaks that dominate their undecided neighbourhood become kept (state 3 = kept this round)
      for (int k = tid; k < m; k += kPPThreads) {
        if (state[k] != 0) continue;
        const int pk = cand[k];
        const double hk = y[pk];
        bool dominated = false;
        for (int j = k - 1; j >= 0 && pk - cand[j] < distance; --j)
          if ((state[j] == 0 || state[j] == 3) && (y[cand[j]] > hk)) { dominated = true; break; }          // equal height: larger index wins
        for (int j = k + 1; !dominated && j < m && cand[j] - pk < distance; ++j)
          if ((state[j] == 0 || state[j] == 3) && (y[cand[j]] >= hk)) dominated = true;
        if (!dominated) state[k] = 3;
      }
      __syncthreads();
      // phase 2: drop undecided peaks within `distance` of a peak kept this round
      for (int k = tid; k < m; k += kPPThreads) {
        if (state[k] != 0) continue;
        const int pk = cand[k];
        bool drop = false;
        for (int j = k - 1; j >= 0 && pk - cand[j] < distance; --j) if (state[j] == 3) { drop = true; break; }
        for (int j = k + 1; !drop && j < m && cand[j] - pk < distance; ++j) if (state[j] == 3) drop = true;
        if (drop) state[k] = 2; else sh_flag = 1;
      }
      __syncthreads();
      for (int k = tid; k < m; k += kPPThreads) if (state[k] == 3) state[k] = 1;
      const int again = sh_flag;
      __syncthreads();
      if (!again) break;
    }
  } else {
    for (int k = tid; k < m; k += kPPThreads) state[k] = 1;
    __syncthreads();
  }

  // ---- E. prominence over the full signal (wlen=None), then the final ordered compaction ----
  for (int k = tid; k < m; k += kPPThreads) {
    if (state[k] != 1) continue;
    const int pk = cand[k];
    const double h = y[pk];
    double lmin = h, rmin = h;
    for (int i = pk; i >= 0 && y[i] <= h; --i) lmin = fmin(lmin, y[i]);
    for (int i = pk; i < T && y[i] <= h; ++i) rmin = fmin(rmin, y[i]);
    if (!(h - fmax(lmin, rmin) >= prominence)) state[k] = 2;
  }
  __syncthreads();
  const int cseg = (m + kPPThreads - 1) / kPPThreads;
  const int k0 = min(m, tid * cseg), k1 = min(m, (tid + 1) * cseg);
  int mine = 0;
  for (int k = k0; k < k1; ++k) mine += (state[k] == 1);
  int total = 0;
  int wr = block_exscan(mine, sh_i, &total);
  for (int k = k0; k < k1; ++k)
    if (state[k] == 1) { if (wr < max_peaks) peaks_out[(size_t)s * max_peaks + wr] = cand[k]; ++wr; }
  if (tid == 0) n_peaks[s] = (cnt > cand_cap) ? 0x7fffffff : total;
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
  const int cand_cap = T / 2 + 1;  // strict local maxima are at least 2 samples apart
  double* y_ws = nullptr;
  int32_t* cand_ws = nullptr;
  int8_t* state_ws = nullptr;
  if (!smoothed) AHA_CUDA(cudaMallocAsync(&y_ws, (size_t)n_streams * T * sizeof(double), st));
  AHA_CUDA(cudaMallocAsync(&cand_ws, (size_t)n_streams * cand_cap * sizeof(int32_t), st));
  AHA_CUDA(cudaMallocAsync(&state_ws, (size_t)n_streams * cand_cap, st));
  find_ticks_kernel<<<n_streams, kPPThreads, 0, st>>>(scores, T, prominence, distance, thresh, peaks, max_peaks, n_peaks,
                                                      smoothed ? smoothed : y_ws, thresh_out, cand_ws, state_ws, cand_cap);
  ++g_aha_launch_count;
  cudaError_t e = cudaGetLastError();
  if (y_ws) cudaFreeAsync(y_ws, st);
  cudaFreeAsync(cand_ws, st);
  cudaFreeAsync(state_ws, st);
  if (e != cudaSuccess) return aha_cuda_fail(e, "find_ticks_kernel", __FILE__, __LINE__);
  return AHA_OK;
}
