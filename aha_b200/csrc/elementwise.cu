// Memory-bound kernels of the ingest-and-score path: coalesced 128-bit HBM access, warp-shuffle reductions.
// Each kernel states the reference op it replaces and its algorithmic bytes.
#include "ops.h"
#include <math.h>

namespace {

constexpr int kNormThreads = 128;
constexpr int kNormMaxVec = 8;  // up to 128 * 8 * 8 = 8192 columns

// RMSNorm / LayerNorm: one 128-thread block per row, the row stays in registers between the statistics
// pass and the normalisation pass (read x once: 2*cols B in, 2*cols B out per row).
// Rounding follows SURVEY.md C.4:  Qwen2RMSNorm  y = bf16(x32 * rsqrt(mean(x32^2) + eps)); out = bf16(w * y)
//                                   nn.LayerNorm  out = bf16((x32 - mean) * rstd * w + b)   (one rounding)
template <bool kLayerNorm>
__global__ void __launch_bounds__(kNormThreads) norm_kernel(const bf16* __restrict__ x, const bf16* __restrict__ w,
                                                            const bf16* __restrict__ b, bf16* __restrict__ out,
                                                            int cols, float eps) {
  pdl_wait();
  pdl_launch_dependents();
  __shared__ float red[32];
  const int row = blockIdx.x;
  const int nvec = cols >> 3;
  const uint4* xr = reinterpret_cast<const uint4*>(x + (size_t)row * cols);
  float v[kNormMaxVec][8];
  float s = 0.f;
#pragma unroll
  for (int it = 0; it < kNormMaxVec; ++it) {
    const int c = threadIdx.x + it * kNormThreads;
    if (c < nvec) {
      const uint4 u = ld_nc_v4(xr + c);
      const float2 a0 = unpack_bf16x2(u.x), a1 = unpack_bf16x2(u.y), a2 = unpack_bf16x2(u.z), a3 = unpack_bf16x2(u.w);
      v[it][0] = a0.x; v[it][1] = a0.y; v[it][2] = a1.x; v[it][3] = a1.y;
      v[it][4] = a2.x; v[it][5] = a2.y; v[it][6] = a3.x; v[it][7] = a3.y;
#pragma unroll
      for (int e = 0; e < 8; ++e) s += kLayerNorm ? v[it][e] : v[it][e] * v[it][e];
    }
  }
  s = block_sum(s, red);
  float mean = 0.f, rstd;
  if (kLayerNorm) {
    mean = s / (float)cols;
    float s2 = 0.f;
#pragma unroll
    for (int it = 0; it < kNormMaxVec; ++it) {
      const int c = threadIdx.x + it * kNormThreads;
      if (c < nvec) {
#pragma unroll
        for (int e = 0; e < 8; ++e) { const float dlt = v[it][e] - mean; s2 += dlt * dlt; }
      }
    }
    s2 = block_sum(s2, red);
    rstd = rsqrtf(s2 / (float)cols + eps);
  } else {
    rstd = rsqrtf(s / (float)cols + eps);
  }
  uint4* orow = reinterpret_cast<uint4*>(out + (size_t)row * cols);
#pragma unroll
  for (int it = 0; it < kNormMaxVec; ++it) {
    const int c = threadIdx.x + it * kNormThreads;
    if (c < nvec) {
      const uint4 wu = *reinterpret_cast<const uint4*>(w + (size_t)c * 8);
      float wv[8];
      { const float2 a0 = unpack_bf16x2(wu.x), a1 = unpack_bf16x2(wu.y), a2 = unpack_bf16x2(wu.z), a3 = unpack_bf16x2(wu.w);
        wv[0] = a0.x; wv[1] = a0.y; wv[2] = a1.x; wv[3] = a1.y; wv[4] = a2.x; wv[5] = a2.y; wv[6] = a3.x; wv[7] = a3.y; }
      float o[8];
      if (kLayerNorm) {
        const uint4 bu = *reinterpret_cast<const uint4*>(b + (size_t)c * 8);
        float bv[8];
        { const float2 a0 = unpack_bf16x2(bu.x), a1 = unpack_bf16x2(bu.y), a2 = unpack_bf16x2(bu.z), a3 = unpack_bf16x2(bu.w);
          bv[0] = a0.x; bv[1] = a0.y; bv[2] = a1.x; bv[3] = a1.y; bv[4] = a2.x; bv[5] = a2.y; bv[6] = a3.x; bv[7] = a3.y; }
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] = (v[it][e] - mean) * rstd * wv[e] + bv[e];
      } else {
#pragma unroll
        for (int e = 0; e < 8; e += 2) {
          float n0 = v[it][e] * rstd, n1 = v[it][e + 1] * rstd;
          bf16r2(n0, n1);                        // the reference casts the normalised row to bf16 before the weight
          o[e] = wv[e] * n0; o[e + 1] = wv[e + 1] * n1;
        }
      }
      uint4 ou;
      ou.x = pack_bf16x2(o[0], o[1]); ou.y = pack_bf16x2(o[2], o[3]);
      ou.z = pack_bf16x2(o[4], o[5]); ou.w = pack_bf16x2(o[6], o[7]);
      st_v4(orow + c, ou);
    }
  }
}

// SigLipImageProcessor.preprocess restricted to already-square frames (siglip_encoder.py:47-67): rescale by 1/255 in
// float64 -> float32, (x - 0.5) / 0.5 in float32, then the caller's .to(bf16).  A pure function of the byte value.
__device__ __forceinline__ float preprocess_value(uint32_t u) {
  const float x = (float)((double)u * (1.0 / 255.0));
  return __fdiv_rn(__fsub_rn(x, 0.5f), 0.5f);
}

__global__ void preprocess_kernel(const uint8_t* __restrict__ in, bf16* __restrict__ out, long long n) {
  __shared__ bf16 lut[256];
  if (threadIdx.x < 256) lut[threadIdx.x] = f2b(preprocess_value(threadIdx.x));
  __syncthreads();
  const long long nvec = n >> 4;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < nvec; i += (long long)gridDim.x * blockDim.x) {
    const uint4 u = ld_nc_v4(in + i * 16);
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
    uint32_t o[8];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const bf16 a = lut[w[k] & 255], b = lut[(w[k] >> 8) & 255], c = lut[(w[k] >> 16) & 255], d = lut[w[k] >> 24];
      o[2 * k] = (uint32_t)__bfloat16_as_ushort(a) | ((uint32_t)__bfloat16_as_ushort(b) << 16);
      o[2 * k + 1] = (uint32_t)__bfloat16_as_ushort(c) | ((uint32_t)__bfloat16_as_ushort(d) << 16);
    }
    st_v4(out + i * 16, make_uint4(o[0], o[1], o[2], o[3]));
    st_v4(out + i * 16 + 8, make_uint4(o[4], o[5], o[6], o[7]));
  }
  if (blockIdx.x == 0) {
    for (long long i = (nvec << 4) + threadIdx.x; i < n; i += blockDim.x) out[i] = lut[in[i]];
  }
}

// Patch extraction for the conv14/14 patch embedding (siglip_encoder.py:156-172) with the preprocess fused in:
// out[(f*side + py)*side + px][c*P*P + ky*P + kx] = pixel(f, c, py*P+ky, px*P+kx); columns >= 3*P*P are zero.
// One thread per (patch row, c, ky): P contiguous input bytes -> P contiguous bf16.
template <bool kU8>
__global__ void patchify_kernel(const uint8_t* __restrict__ fu8, const bf16* __restrict__ fpx, bf16* __restrict__ out,
                                int n, int image, int patch, int kpad) {
  pdl_wait();
  pdl_launch_dependents();
  __shared__ bf16 lut[256];
  if (kU8) {
    if (threadIdx.x < 256) lut[threadIdx.x] = f2b(preprocess_value(threadIdx.x));
    __syncthreads();
  }
  const int side = image / patch;
  const int kreal = 3 * patch * patch;
  const int segs = 3 * patch + 1;  // (c, ky) pairs + one zero-pad segment
  const long long total = (long long)n * side * side * segs;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int seg = (int)(idx % segs);
    const long long row = idx / segs;
    bf16* o = out + row * kpad;
    if (seg == segs - 1) {
      for (int k = kreal; k < kpad; ++k) o[k] = f2b(0.f);
      continue;
    }
    const int c = seg / patch, ky = seg % patch;
    const int px = (int)(row % side);
    const int py = (int)((row / side) % side);
    const int f = (int)(row / ((long long)side * side));
    const size_t src = (((size_t)f * 3 + c) * image + (size_t)py * patch + ky) * image + (size_t)px * patch;
    o += (c * patch + ky) * patch;
    if (kU8) {
      for (int kx = 0; kx < patch; ++kx) o[kx] = lut[fu8[src + kx]];
    } else {
      for (int kx = 0; kx < patch; ++kx) o[kx] = fpx[src + kx];
    }
  }
}

// post_projector_pooling (video_head_live_llava_qwen.py:117-136).  x: [F, side*side, C] -> out [F, os*os, C].
// bilinear: F.interpolate(align_corners=False): src = scale*(dst+0.5)-0.5 clamped at 0, 4 taps, fp32 opmath, one rounding
// average : avg_pool2d(stride)   max: max_pool2d(stride)   adaptive: adaptive_avg_pool2d (models/vision_live.py:21-26)
// One thread per (frame, output token, 8 channels): 4..25 x 16-byte loads, one 16-byte store.
__global__ void pool_kernel(const bf16* __restrict__ x, bf16* __restrict__ out, int frames, int side, int channels,
                            int stride, int mode, int os) {
  pdl_wait();
  pdl_launch_dependents();
  const int cvec = channels >> 3;
  const long long total = (long long)frames * os * os * cvec;
  const float scale = (float)side / (float)os;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int cv = (int)(idx % cvec);
    const long long t = idx / cvec;
    const int ox = (int)(t % os), oy = (int)((t / os) % os);
    const int f = (int)(t / ((long long)os * os));
    const bf16* base = x + (size_t)f * side * side * channels + (size_t)cv * 8;
    float acc[8];
    auto load8 = [&](int yy, int xx, float (&r)[8]) {
      const uint4 u = *reinterpret_cast<const uint4*>(base + ((size_t)yy * side + xx) * channels);
      const float2 a0 = unpack_bf16x2(u.x), a1 = unpack_bf16x2(u.y), a2 = unpack_bf16x2(u.z), a3 = unpack_bf16x2(u.w);
      r[0] = a0.x; r[1] = a0.y; r[2] = a1.x; r[3] = a1.y; r[4] = a2.x; r[5] = a2.y; r[6] = a3.x; r[7] = a3.y;
    };
    if (mode == AHA_POOL_BILINEAR) {
      float sy = scale * ((float)oy + 0.5f) - 0.5f; if (sy < 0.f) sy = 0.f;
      float sx = scale * ((float)ox + 0.5f) - 0.5f; if (sx < 0.f) sx = 0.f;
      const int y0 = (int)sy, x0 = (int)sx;
      const int y1 = y0 + (y0 < side - 1 ? 1 : 0), x1 = x0 + (x0 < side - 1 ? 1 : 0);
      const float ly1 = sy - (float)y0, lx1 = sx - (float)x0;
      const float ly0 = 1.f - ly1, lx0 = 1.f - lx1;
      float a[8], b[8], c[8], d[8];
      load8(y0, x0, a); load8(y0, x1, b); load8(y1, x0, c); load8(y1, x1, d);
#pragma unroll
      for (int e = 0; e < 8; ++e)
        acc[e] = __fadd_rn(__fmul_rn(ly0, __fadd_rn(__fmul_rn(lx0, a[e]), __fmul_rn(lx1, b[e]))),
                           __fmul_rn(ly1, __fadd_rn(__fmul_rn(lx0, c[e]), __fmul_rn(lx1, d[e]))));
    } else {
      int ys, ye, xs, xe;
      if (mode == AHA_POOL_ADAPTIVE) {
        ys = (oy * side) / os; ye = ((oy + 1) * side + os - 1) / os;
        xs = (ox * side) / os; xe = ((ox + 1) * side + os - 1) / os;
      } else {
        ys = oy * stride; ye = ys + stride; xs = ox * stride; xe = xs + stride;
      }
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[e] = (mode == AHA_POOL_MAX) ? -INFINITY : 0.f;
      for (int yy = ys; yy < ye; ++yy)
        for (int xx = xs; xx < xe; ++xx) {
          float r[8];
          load8(yy, xx, r);
#pragma unroll
          for (int e = 0; e < 8; ++e) acc[e] = (mode == AHA_POOL_MAX) ? fmaxf(acc[e], r[e]) : acc[e] + r[e];
        }
      if (mode != AHA_POOL_MAX) {
        const float cnt = (float)((ye - ys) * (xe - xs));
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] = acc[e] / cnt;
      }
    }
    uint4 ou;
    ou.x = pack_bf16x2(acc[0], acc[1]); ou.y = pack_bf16x2(acc[2], acc[3]);
    ou.z = pack_bf16x2(acc[4], acc[5]); ou.w = pack_bf16x2(acc[6], acc[7]);
    st_v4(out + ((size_t)t * channels + (size_t)cv * 8), ou);
  }
}

// Qwen2RotaryEmbedding (transformers 4.49): freqs = pos * inv_freq in fp32, emb = cat(freqs, freqs),
// cos/sin in fp32, cast to bf16.  inv_freq is computed on the host exactly as torch does (model.cu).
__global__ void rope_table_kernel(RopeTableArgs a, int q, int d, const float* __restrict__ inv_freq,
                                  bf16* __restrict__ cos_out, bf16* __restrict__ sin_out) {
  pdl_wait();
  pdl_launch_dependents();
  const int s = blockIdx.y;
  const int half = d >> 1;
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < q * half; idx += gridDim.x * blockDim.x) {
    const int i = idx / half, j = idx % half;
    const float fr = __fmul_rn((float)(a.pos0[s] + i), inv_freq[j]);
    const bf16 c = f2b(cosf(fr)), sn = f2b(sinf(fr));
    const size_t o = ((size_t)s * q + i) * d + j;
    cos_out[o] = c; cos_out[o + half] = c;
    sin_out[o] = sn; sin_out[o + half] = sn;
  }
}

// torch.cat([embed_tokens(ids), frame_embeds]) (test/inference.py:211-214): one block per output row.
__global__ void embed_assemble_kernel(const bf16* __restrict__ table, const int64_t* __restrict__ ids, int n_ids,
                                      const bf16* __restrict__ frame_embeds, bf16* __restrict__ out, int hidden,
                                      int vocab) {
  pdl_wait();
  pdl_launch_dependents();
  const int row = blockIdx.x;
  const bf16* src;
  if (row < n_ids) {
    long long id = ids[row];
    if (id < 0 || id >= vocab) id = 0;  // memory safety only: torch would raise, and so do the Python wrappers (model.py:
                                        // assemble_chunk and get_input_embeddings check the range before this launch)
    src = table + (size_t)id * hidden;
  } else {
    src = frame_embeds + (size_t)(row - n_ids) * hidden;
  }
  const uint4* s4 = reinterpret_cast<const uint4*>(src);
  uint4* o4 = reinterpret_cast<uint4*>(out + (size_t)row * hidden);
  for (int c = threadIdx.x; c < (hidden >> 3); c += blockDim.x) o4[c] = s4[c];
}

// informative_head / relevance_head / uncertainty_head on every row (video_head_live_llava_qwen.py:185-188):
// bf16 Linear (fp32 accumulate, one rounding to bf16) -> .float() -> sigmoid for relevance; and the host-side readout of
// test/inference.py:222-227 for the last row of every stream: softmax(inf)[1], rel, exp(unc).
__global__ void __launch_bounds__(128) heads_kernel(const bf16* __restrict__ h, const bf16* __restrict__ w_inf,
                                                    const bf16* __restrict__ w_rel, const bf16* __restrict__ w_unc,
                                                    int hidden, int q, float* __restrict__ informative,
                                                    float* __restrict__ relevance, float* __restrict__ uncertainty,
                                                    float* __restrict__ scores3) {
  pdl_wait();
  pdl_launch_dependents();
  __shared__ float red[4][4];
  const int row = blockIdx.x;
  const uint4* hr = reinterpret_cast<const uint4*>(h + (size_t)row * hidden);
  const uint4* w0 = reinterpret_cast<const uint4*>(w_inf);
  const uint4* w1 = reinterpret_cast<const uint4*>(w_inf + hidden);
  const uint4* w2 = reinterpret_cast<const uint4*>(w_rel);
  const uint4* w3 = reinterpret_cast<const uint4*>(w_unc);
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  auto dot8 = [](const uint4& a, const uint4& b) {
    const float2 a0 = unpack_bf16x2(a.x), a1 = unpack_bf16x2(a.y), a2 = unpack_bf16x2(a.z), a3 = unpack_bf16x2(a.w);
    const float2 b0 = unpack_bf16x2(b.x), b1 = unpack_bf16x2(b.y), b2 = unpack_bf16x2(b.z), b3 = unpack_bf16x2(b.w);
    return a0.x * b0.x + a0.y * b0.y + a1.x * b1.x + a1.y * b1.y + a2.x * b2.x + a2.y * b2.y + a3.x * b3.x + a3.y * b3.y;
  };
  for (int c = threadIdx.x; c < (hidden >> 3); c += blockDim.x) {
    const uint4 x = hr[c];
    acc[0] += dot8(x, w0[c]); acc[1] += dot8(x, w1[c]); acc[2] += dot8(x, w2[c]); acc[3] += dot8(x, w3[c]);
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const float s = warp_sum(acc[k]);
    if (lane == 0) red[k][warp] = s;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    float r[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) r[k] = bf16r(red[k][0] + red[k][1] + red[k][2] + red[k][3]);
    const float rel = 1.f / (1.f + expf(-r[2]));
    if (informative) { informative[(size_t)row * 2] = r[0]; informative[(size_t)row * 2 + 1] = r[1]; }
    if (relevance) relevance[row] = rel;
    if (uncertainty) uncertainty[row] = r[3];
    if (scores3 && (row % q) == q - 1) {
      const int s = row / q;
      const float m = fmaxf(r[0], r[1]);
      const float e0 = expf(r[0] - m), e1 = expf(r[1] - m);
      scores3[s * 3 + 0] = e1 / (e0 + e1);
      scores3[s * 3 + 1] = rel;
      scores3[s * 3 + 2] = expf(r[3]);
    }
  }
}

__global__ void copy_rows_kernel(const bf16* __restrict__ src, int src_ld, bf16* __restrict__ dst, int dst_ld, int rows,
                                 int cols, int dst_row0, int dst_row_step) {
  const long long total = (long long)rows * cols;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const int r = (int)(idx / cols), c = (int)(idx % cols);
    dst[(size_t)(dst_row0 + r * dst_row_step) * dst_ld + c] = src[(size_t)r * src_ld + c];
  }
}

inline int grid_for(long long work_items, int threads, int cap = 148 * 16) {
  long long g = (work_items + threads - 1) / threads;
  if (g < 1) g = 1;
  if (g > cap) g = cap;
  return (int)g;
}

}  // namespace

// nn.LayerNorm for many short rows (the SigLIP tower: 23 328 rows x 1152): one WARP per row, the row lives in registers
// (kV 16-byte loads per lane, all in flight at once), statistics by shuffles only - no block barrier, 8 rows per CTA.
// Same formula and roundings as norm_kernel<true> (two-pass variance, one bf16 rounding at the end).
template <int kV>
__global__ void __launch_bounds__(256) layernorm_warp_kernel(const bf16* __restrict__ x, const bf16* __restrict__ w,
                                                             const bf16* __restrict__ b, bf16* __restrict__ out, int rows,
                                                             int cols, float eps) {
  pdl_wait();
  pdl_launch_dependents();
  const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (row >= rows) return;
  const int nvec = cols >> 3;
  const uint4* xr = reinterpret_cast<const uint4*>(x + (size_t)row * cols);
  uint4 u[kV];
#pragma unroll
  for (int it = 0; it < kV; ++it) {
    const int c = lane + it * 32;
    u[it] = c < nvec ? ld_nc_v4(xr + c) : make_uint4(0, 0, 0, 0);
  }
  float v[kV][8];
  float s = 0.f;
#pragma unroll
  for (int it = 0; it < kV; ++it) {
    const float2 a0 = unpack_bf16x2(u[it].x), a1 = unpack_bf16x2(u[it].y), a2 = unpack_bf16x2(u[it].z), a3 = unpack_bf16x2(u[it].w);
    v[it][0] = a0.x; v[it][1] = a0.y; v[it][2] = a1.x; v[it][3] = a1.y;
    v[it][4] = a2.x; v[it][5] = a2.y; v[it][6] = a3.x; v[it][7] = a3.y;
#pragma unroll
    for (int e = 0; e < 8; ++e) s += v[it][e];          // padding lanes hold zeros
  }
  const float mean = warp_sum(s) / (float)cols;
  float s2 = 0.f;
#pragma unroll
  for (int it = 0; it < kV; ++it) {
    if (lane + it * 32 < nvec) {
#pragma unroll
      for (int e = 0; e < 8; ++e) { const float dlt = v[it][e] - mean; s2 += dlt * dlt; }
    }
  }
  const float rstd = rsqrtf(warp_sum(s2) / (float)cols + eps);
  uint4* orow = reinterpret_cast<uint4*>(out + (size_t)row * cols);
#pragma unroll
  for (int it = 0; it < kV; ++it) {
    const int c = lane + it * 32;
    if (c < nvec) {
      const uint4 wu = *reinterpret_cast<const uint4*>(w + (size_t)c * 8);
      const uint4 bu = *reinterpret_cast<const uint4*>(b + (size_t)c * 8);
      const uint32_t ww[4] = {wu.x, wu.y, wu.z, wu.w}, bw[4] = {bu.x, bu.y, bu.z, bu.w};
      uint32_t o[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float2 wf = unpack_bf16x2(ww[e]), bf = unpack_bf16x2(bw[e]);
        o[e] = pack_bf16x2((v[it][2 * e] - mean) * rstd * wf.x + bf.x, (v[it][2 * e + 1] - mean) * rstd * wf.y + bf.y);
      }
      st_v4(orow + c, make_uint4(o[0], o[1], o[2], o[3]));
    }
  }
}

// ---- folded RMSNorm helpers (see GemmCall in gemm.h) ----
// row copy + sum of squares of the copied row (the residual stream entering layer 0): block per row
__global__ void __launch_bounds__(256) copy_sumsq_kernel(const bf16* __restrict__ x, bf16* __restrict__ out,
                                                         float* __restrict__ sumsq, int cols) {
  pdl_wait();
  pdl_launch_dependents();
  __shared__ float red[8];
  const uint4* src = reinterpret_cast<const uint4*>(x + (size_t)blockIdx.x * cols);
  uint4* dst = reinterpret_cast<uint4*>(out + (size_t)blockIdx.x * cols);
  float s = 0.f;
  for (int c = threadIdx.x; c < (cols >> 3); c += blockDim.x) {
    const uint4 u = src[c];
    dst[c] = u;
    const float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), d = unpack_bf16x2(u.z), e = unpack_bf16x2(u.w);
    s += a.x * a.x + a.y * a.y + b.x * b.x + b.y * b.y + d.x * d.x + d.y * d.y + e.x * e.x + e.y * e.y;
  }
  s = warp_sum(s);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += red[w];
    sumsq[blockIdx.x] = t;
  }
}

// W[r][c] = bf16(W[r][c] * v[c]): the norm weight folded into the columns of the matrix that consumes the normalised row
__global__ void fold_columns_kernel(bf16* __restrict__ w, const bf16* __restrict__ v, long long rows, int cols) {
  const long long n = rows * (long long)(cols >> 3);
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % (cols >> 3));
    uint4 u = reinterpret_cast<uint4*>(w)[i];
    const uint4 s = reinterpret_cast<const uint4*>(v)[c];
    uint32_t a[4] = {u.x, u.y, u.z, u.w};
    const uint32_t b[4] = {s.x, s.y, s.z, s.w};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float2 x = unpack_bf16x2(a[e]), y = unpack_bf16x2(b[e]);
      a[e] = pack_bf16x2(x.x * y.x, x.y * y.y);
    }
    reinterpret_cast<uint4*>(w)[i] = make_uint4(a[0], a[1], a[2], a[3]);
  }
}

int op_copy_sumsq(const bf16* x, bf16* out, float* sumsq, int rows, int cols, cudaStream_t st) {
  AHA_REQUIRE(rows > 0 && cols > 0 && cols % 8 == 0, "copy_sumsq: bad shape");
  return launch_pdl(copy_sumsq_kernel, dim3(rows), dim3(256), 0, st, x, out, sumsq, cols);
}

int op_fold_columns(bf16* w, const bf16* v, long long rows, int cols, cudaStream_t st) {
  AHA_REQUIRE(rows > 0 && cols > 0 && cols % 8 == 0, "fold_columns: bad shape");
  const long long n = rows * (long long)(cols >> 3);
  const int blocks = (int)std::min<long long>((n + 255) / 256, 148 * 8);
  fold_columns_kernel<<<blocks, 256, 0, st>>>(w, v, rows, cols);
  AHA_CHECK_LAUNCH();
  return AHA_OK;
}

int op_rmsnorm(const bf16* x, const bf16* w, bf16* out, int rows, int cols, float eps, cudaStream_t st) {
  AHA_REQUIRE(rows > 0 && cols > 0 && cols % 8 == 0 && cols <= kNormThreads * kNormMaxVec * 8, "rmsnorm: bad shape");
  return launch_pdl(norm_kernel<false>, dim3(rows), dim3(kNormThreads), 0, st, x, w, (const bf16*)nullptr, out, cols, eps);
}

int op_layernorm(const bf16* x, const bf16* w, const bf16* b, bf16* out, int rows, int cols, float eps, cudaStream_t st) {
  AHA_REQUIRE(rows > 0 && cols > 0 && cols % 8 == 0 && cols <= kNormThreads * kNormMaxVec * 8, "layernorm: bad shape");
  if (rows >= 1024 && cols <= 32 * 5 * 8)   // many short rows: warp per row
    return launch_pdl(layernorm_warp_kernel<5>, dim3((rows + 7) / 8), dim3(256), 0, st, x, w, b, out, rows, cols, eps);
  return launch_pdl(norm_kernel<true>, dim3(rows), dim3(kNormThreads), 0, st, x, w, b, out, cols, eps);
}

int op_preprocess(const uint8_t* in, bf16* out, long long n, cudaStream_t st) {
  AHA_REQUIRE(n >= 0, "preprocess: negative size");
  if (n == 0) return AHA_OK;
  AHA_REQUIRE((reinterpret_cast<uintptr_t>(in) & 15) == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0,
              "preprocess: pointers must be 16-byte aligned");
  preprocess_kernel<<<grid_for(n / 16 + 1, 256), 256, 0, st>>>(in, out, n);
  ++g_aha_launch_count;
  AHA_CHECK_LAUNCH();
  return AHA_OK;
}

int op_patchify(const uint8_t* frames_u8, const bf16* pixels, bf16* out, int n, int image, int patch, int kpad,
                cudaStream_t st) {
  AHA_REQUIRE((frames_u8 != nullptr) != (pixels != nullptr), "patchify: give exactly one of frames_u8 / pixel_values");
  AHA_REQUIRE(n > 0 && image >= patch && kpad >= 3 * patch * patch, "patchify: bad geometry");
  const int side = image / patch;
  const long long total = (long long)n * side * side * (3 * patch + 1);
  if (frames_u8) return launch_pdl(patchify_kernel<true>, dim3(grid_for(total, 256, 148 * 32)), dim3(256), 0, st, frames_u8, (const bf16*)nullptr, out, n, image, patch, kpad);
  return launch_pdl(patchify_kernel<false>, dim3(grid_for(total, 256, 148 * 32)), dim3(256), 0, st, (const uint8_t*)nullptr, pixels, out, n, image, patch, kpad);
}

int pool_out_side(int side, int stride, int mode) {
  if (mode == AHA_POOL_BILINEAR) return (side + stride - 1) / stride;
  if (mode == AHA_POOL_ADAPTIVE) return stride;  // `stride` carries the target side (frame_token_pooled)
  return side / stride;
}

int op_pool(const bf16* x, bf16* out, int frames, int side, int channels, int stride, int mode, cudaStream_t st) {
  AHA_REQUIRE(mode >= 0 && mode <= 3, "Unexpected mm_spatial_pool_mode");
  AHA_REQUIRE(frames > 0 && side > 0 && channels % 8 == 0 && stride > 0, "pool: bad shape");
  const int os = pool_out_side(side, stride, mode);
  AHA_REQUIRE(os > 0, "pool: empty output");
  const long long total = (long long)frames * os * os * (channels / 8);
  return launch_pdl(pool_kernel, dim3(grid_for(total, 256)), dim3(256), 0, st, x, out, frames, side, channels, stride, mode, os);
}

int op_rope_table(const RopeTableArgs& a, int n_streams, int q, int d, const float* inv_freq_dev, bf16* cos_out,
                  bf16* sin_out, cudaStream_t st) {
  AHA_REQUIRE(n_streams > 0 && n_streams <= 32 && q > 0 && d % 2 == 0, "rope_table: bad shape");
  dim3 grid((q * (d / 2) + 127) / 128, n_streams);
  return launch_pdl(rope_table_kernel, grid, dim3(128), 0, st, a, q, d, inv_freq_dev, cos_out, sin_out);
}

int op_embed_assemble(const bf16* table, const int64_t* ids, int n_ids, const bf16* frame_embeds, int n_frame_tokens,
                      bf16* out, int hidden, int vocab, cudaStream_t st) {
  AHA_REQUIRE(n_ids >= 0 && n_frame_tokens >= 0 && hidden % 8 == 0, "assemble: bad shape");
  if (n_ids + n_frame_tokens == 0) return AHA_OK;
  return launch_pdl(embed_assemble_kernel, dim3(n_ids + n_frame_tokens), dim3(128), 0, st, table, ids, n_ids, frame_embeds, out, hidden, vocab);
}

int op_heads(const bf16* h, const bf16* w_inf, const bf16* w_rel, const bf16* w_unc, int rows, int hidden, int q,
             float* informative, float* relevance, float* uncertainty, float* scores3, cudaStream_t st) {
  AHA_REQUIRE(rows > 0 && hidden % 8 == 0 && q > 0, "heads: bad shape");
  return launch_pdl(heads_kernel, dim3(rows), dim3(128), 0, st, h, w_inf, w_rel, w_unc, hidden, q, informative, relevance, uncertainty, scores3);
}

int op_copy_rows(const bf16* src, int src_ld, bf16* dst, int dst_ld, int rows, int cols, int dst_row0, int dst_row_step,
                 cudaStream_t st) {
  if (rows <= 0 || cols <= 0) return AHA_OK;
  copy_rows_kernel<<<grid_for((long long)rows * cols, 256, 148 * 32), 256, 0, st>>>(src, src_ld, dst, dst_ld, rows, cols,
                                                                                  dst_row0, dst_row_step);
  ++g_aha_launch_count;
  AHA_CHECK_LAUNCH();
  return AHA_OK;
}

int op_fill_zero(void* dst, size_t bytes, cudaStream_t st) {
  AHA_CUDA(cudaMemsetAsync(dst, 0, bytes, st));
  return AHA_OK;
}
