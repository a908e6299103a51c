// Score fusion and highlight selection on the per-frame score stream, float64 like the Python
// floats the reference computes with, every multiply and add rounded separately (no FMA contraction) in the reference's
// evaluation order so that results are bit-identical:
//   fuse_scores        test/grid_search.py:58-65 (and :95-104): alpha*inf + beta*rel, minus (unc - thr)*eps when unc >= thr
//   knapsack_select    test/highlight_generator.py:8-37: 0/1 knapsack, every frame costs 1, value = inf*alpha + rel*beta +
//                      unc*epsilon; the DP recurrence and the backtracking rule `dp[i][c] != dp[i-1][c]` are evaluated
//                      exactly (one warp per stream: lane j owns the budgets j + 32 w, W 32-bit decision words per frame
//                      replace the (n+1) x (budget+1) table; any budget up to 1023)
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

// one warp per stream; lane j holds dp[i][b] for the budgets b = j + 32 w, w < W (W decision words per frame: bit j of word w
// says dp[i][j + 32 w] != dp[i-1][j + 32 w]).  dp[i-1][b-1] sits in the lane below, or - for lane 0 - in lane 31 of word w - 1.
template <int W>
__global__ void knapsack_kernel(const double* __restrict__ inf, const double* __restrict__ rel,
                                const double* __restrict__ unc, int n_streams, int T, int budget, double alpha, double beta,
                                double eps, uint32_t* __restrict__ bits, uint8_t* __restrict__ selected) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= n_streams) return;
  const size_t base = (size_t)warp * T;
  double dp[W];
#pragma unroll
  for (int w = 0; w < W; ++w) dp[w] = 0.0;   // dp[0][b] = 0
  for (int i = 0; i < T; ++i) {
    const double value = __dadd_rn(__dadd_rn(__dmul_rn(inf[base + i], alpha), __dmul_rn(rel[base + i], beta)),
                                   __dmul_rn(unc[base + i], eps));
    double carry = 0.0;                                                  // dp[i-1][32 w - 1] for lane 0
#pragma unroll
    for (int w = 0; w < W; ++w) {
      const double up = __shfl_up_sync(0xffffffffu, dp[w], 1);
      const double top = __shfl_sync(0xffffffffu, dp[w], 31);
      const double left = lane == 0 ? carry : up;                        // dp[i-1][b-1]
      const int bud = lane + 32 * w;
      double nd = dp[w];
      if (bud >= 1 && bud <= budget) nd = fmax(dp[w], __dadd_rn(left, value));   // cost 1 <= b
      const uint32_t word = __ballot_sync(0xffffffffu, nd != dp[w]);     // bit j: dp[i][b] != dp[i-1][b]
      if (lane == 0) bits[(base + i) * W + w] = word;
      carry = top;
      dp[w] = nd;
    }
  }
  __syncwarp();
  // backtracking (sequential, lane 0); the other lanes clear the output first
  for (int i = lane; i < T; i += 32) selected[base + i] = 0;
  __syncwarp();
  if (lane == 0) {
    int rc = budget;
    for (int i = T - 1; i >= 0; --i) {
      if ((bits[(base + i) * W + (rc >> 5)] >> (rc & 31)) & 1u) { selected[base + i] = 1; --rc; }
    }
  }
}

template <int W>
void launch_knapsack(const double* inf, const double* rel, const double* unc, int n_streams, int T, int budget, double alpha,
                     double beta, double eps, uint32_t* bits, uint8_t* selected, cudaStream_t st) {
  const int warps_per_block = 4;
  knapsack_kernel<W><<<(n_streams + warps_per_block - 1) / warps_per_block, warps_per_block * 32, 0, st>>>(
      inf, rel, unc, n_streams, T, budget, alpha, beta, eps, bits, selected);
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

int aha_knapsack_words(int max_duration) {
  int w = 1;
  while (w * 32 < max_duration + 1 && w < 32) w *= 2;
  return w;
}

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
  AHA_REQUIRE(max_duration >= 0 && max_duration <= 1023, "aha_knapsack_select: max_duration must be in [0, 1023]");
  // W = decision words per frame (aha_knapsack_words): the smallest of 1, 2, 4, 8, 16, 32 that holds max_duration + 1 bits
  cudaStream_t st = (cudaStream_t)stream;
  switch (aha_knapsack_words(max_duration)) {
    case 1: launch_knapsack<1>(inf, rel, unc, n_streams, T, max_duration, alpha, beta, epsilon, scratch_bits, selected, st); break;
    case 2: launch_knapsack<2>(inf, rel, unc, n_streams, T, max_duration, alpha, beta, epsilon, scratch_bits, selected, st); break;
    case 4: launch_knapsack<4>(inf, rel, unc, n_streams, T, max_duration, alpha, beta, epsilon, scratch_bits, selected, st); break;
    case 8: launch_knapsack<8>(inf, rel, unc, n_streams, T, max_duration, alpha, beta, epsilon, scratch_bits, selected, st); break;
    case 16: launch_knapsack<16>(inf, rel, unc, n_streams, T, max_duration, alpha, beta, epsilon, scratch_bits, selected, st); break;
    default: launch_knapsack<32>(inf, rel, unc, n_streams, T, max_duration, alpha, beta, epsilon, scratch_bits, selected, st); break;
  }
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
