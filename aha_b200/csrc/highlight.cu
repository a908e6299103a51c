// Score fusion and highlight selection on the per-frame score stream, float64 like the Python
// floats the reference computes with, every multiply and add rounded separately (no FMA contraction) in the reference's
// evaluation order so that results are bit-identical:
//   fuse_scores        test/grid_search.py:58-65 (and :95-104): alpha*inf + beta*rel, minus (unc - thr)*eps when unc >= thr
//   knapsack_select    test/highlight_generator.py:8-37: 0/1 knapsack, every frame costs 1, value = inf*alpha + rel*beta +
//                      unc*epsilon; the DP recurrence and the backtracking rule `dp[i][c] != dp[i-1][c]` are evaluated
//                      exactly (one warp per stream: lane j owns budget j, a 32-bit decision word per frame replaces the
//                      (n+1) x (budget+1) table)
//   highlight_mask     test/highlight_generator.py:76-86: every selected frame widened to [gt - half, gt + half] clipped
//                      to [0, last_frame]
#include "ops.h"
#include <algorithm>

namespace {

__global__ void fuse_scores_kernel(const double* __restrict__ inf, const double* __restrict__ rel,
                                   const double* __restrict__ unc, long long n, double alpha, double beta, double eps,
                                   double thr, double* __restrict__ out) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    double s = __dadd_rn(__dmul_rn(alpha, inf[i]), __dmul_rn(beta, rel[i]));
    const double u = unc[i];
    if (u >= thr) s = __dsub_rn(s, __dmul_rn(__dsub_rn(u, thr), eps));
    out[i] = s;
  }
}

// one warp per stream; lane j holds dp[i][j] for budget j (j <= budget <= 31)
__global__ void knapsack_kernel(const double* __restrict__ inf, const double* __restrict__ rel,
                                const double* __restrict__ unc, int n_streams, int T, int budget, double alpha, double beta,
                                double eps, uint32_t* __restrict__ bits, uint8_t* __restrict__ selected) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= n_streams) return;
  const size_t base = (size_t)warp * T;
  double dp = 0.0;   // dp[0][j] = 0
  for (int i = 0; i < T; ++i) {
    const double value = __dadd_rn(__dadd_rn(__dmul_rn(inf[base + i], alpha), __dmul_rn(rel[base + i], beta)),
                                   __dmul_rn(unc[base + i], eps));
    const double left = __shfl_up_sync(0xffffffffu, dp, 1);            // dp[i-1][j-1]
    double nd = dp;
    if (lane >= 1 && lane <= budget) nd = fmax(dp, __dadd_rn(left, value));   // cost 1 <= j
    const uint32_t word = __ballot_sync(0xffffffffu, nd != dp);        // bit j: dp[i][j] != dp[i-1][j]
    if (lane == 0) bits[base + i] = word;
    dp = nd;
  }
  __syncwarp();
  // backtracking (sequential, lane 0); the other lanes clear the output first
  for (int i = lane; i < T; i += 32) selected[base + i] = 0;
  __syncwarp();
  if (lane == 0) {
    int rc = budget;
    for (int i = T - 1; i >= 0; --i) {
      if ((bits[base + i] >> rc) & 1u) { selected[base + i] = 1; --rc; }
    }
  }
}

__global__ void highlight_mask_kernel(const uint8_t* __restrict__ selected, const int32_t* __restrict__ gt_frames, int T,
                                      int half, int last_frame, uint8_t* __restrict__ mask) {
  const int s = blockIdx.y;
  const int len = last_frame + 1;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < T; i += gridDim.x * blockDim.x) {
    if (!selected[(size_t)s * T + i]) continue;
    const int g = gt_frames[i];
    const int a = max(0, g - half), b = min(len, g + half + 1);
    for (int f = a; f < b; ++f) mask[(size_t)s * len + f] = 1;   // benign race: every writer stores 1
  }
}

}  // namespace

extern "C" {

int aha_fuse_scores(const double* inf, const double* rel, const double* unc, long long n, double alpha, double beta,
                    double epsilon, double uncertainty_threshold, double* out, void* stream) {
  AHA_REQUIRE(inf && rel && unc && out && n > 0, "aha_fuse_scores: bad arguments");
  const int blocks = (int)std::min<long long>((n + 255) / 256, 148 * 16);
  fuse_scores_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(inf, rel, unc, n, alpha, beta, epsilon,
                                                               uncertainty_threshold, out);
  ++g_aha_launch_count;
  AHA_CHECK_LAUNCH();
  return AHA_OK;
}

int aha_knapsack_select(const double* inf, const double* rel, const double* unc, int n_streams, int T, int max_duration,
                        double alpha, double beta, double epsilon, uint32_t* scratch_bits, uint8_t* selected,
                        void* stream) {
  AHA_REQUIRE(inf && rel && unc && scratch_bits && selected && n_streams > 0 && T > 0, "aha_knapsack_select: bad arguments");
  AHA_REQUIRE(max_duration >= 0 && max_duration <= 31, "aha_knapsack_select: max_duration must be in [0, 31]");
  const int warps_per_block = 4;
  knapsack_kernel<<<(n_streams + warps_per_block - 1) / warps_per_block, warps_per_block * 32, 0, (cudaStream_t)stream>>>(
      inf, rel, unc, n_streams, T, max_duration, alpha, beta, epsilon, scratch_bits, selected);
  ++g_aha_launch_count;
  AHA_CHECK_LAUNCH();
  return AHA_OK;
}

int aha_highlight_mask(const uint8_t* selected, const int32_t* gt_frames, int n_streams, int T, int half_width,
                       int last_frame, uint8_t* mask, void* stream) {
  AHA_REQUIRE(selected && gt_frames && mask && n_streams > 0 && T > 0 && half_width >= 0 && last_frame >= 0,
              "aha_highlight_mask: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  AHA_CUDA(cudaMemsetAsync(mask, 0, (size_t)n_streams * (last_frame + 1), st));
  dim3 grid((T + 255) / 256, n_streams);
  highlight_mask_kernel<<<grid, 256, 0, st>>>(selected, gt_frames, T, half_width, last_frame, mask);
  ++g_aha_launch_count;
  AHA_CHECK_LAUNCH();
  return AHA_OK;
}

}  // extern "C"
