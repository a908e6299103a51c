// Frame ingest on the device: decoded frames [n, h, w, 3] uint8 -> the reference's
// [n, 3, 384, 384] uint8 canvas (aspect-preserving resize, centred on black, CHW), bit-identical to
//   flavour 0: test/inference.py:540-563          cv2.resize (8-bit INTER_LINEAR) + copyMakeBorder + COLOR_BGR2RGB
//   flavour 1: test/live_infer_for_video.py:107-119  PIL Image.resize (BICUBIC, 8 bits per channel) + ImageOps.expand
// The resampling arithmetic is OpenCV's / Pillow's (third-party, not under /root/reference).  Their coefficient tables
// are functions of the geometry only: they are computed on the host in the libraries' own float / double evaluation
// order (resize.cpp resizeGeneric_ setup; Resample.c precompute_coeffs + normalize_coeffs_8bpc), cached per geometry and
// uploaded once; the kernels do the integer arithmetic.  HBM-bound byte work: every canvas byte is written once, by
// consecutive threads along x of one colour plane; the source is read through L1/L2 only where taps land.
#include "ops.h"
#include <algorithm>
#include <cmath>
#include <mutex>
#include <vector>

namespace {

constexpr int kPilBits = 22;     // Resample.c PRECISION_BITS = 32 - 8 - 2

struct IngestTables {            // device arrays of one (h, w, flavour, out) geometry
  int h = 0, w = 0, flavour = -1, out = 0, device = -1;
  int new_w = 0, new_h = 0, left = 0, top = 0, area_fast = 0, ksx = 0, ksy = 0;
  int th = 1, tile_rows = 0;      // PIL flavour: canvas rows per CTA and the most source rows under one tile
  int* x_lo = nullptr;  int* x_cnt = nullptr;  int* x_k = nullptr;   // [new_w], [new_w], [new_w, ksx]
  int* y_lo = nullptr;  int* y_cnt = nullptr;  int* y_k = nullptr;   // cv2: y_lo = row0, y_cnt = row1
  void* block = nullptr;
};
// tables of the last few geometries (a demo alternates between a video geometry and single pictures); the oldest entry
// is freed only after a device-wide synchronise, whatever streams its launches ran on
constexpr size_t kTableCacheEntries = 8;
std::vector<IngestTables> g_tabs;
IngestTables g_tab;               // the entry of the current call
std::mutex g_tab_mutex;

// ---- OpenCV 8-bit INTER_LINEAR: index + two 11-bit coefficients per destination coordinate
void cv2_axis(int src, int dst, bool horizontal, std::vector<int>& lo, std::vector<int>& hi, std::vector<int>& k) {
  const double scale = 1.0 / ((double)dst / (double)src);
  lo.resize(dst); hi.resize(dst); k.resize(2 * (size_t)dst);
  for (int d = 0; d < dst; ++d) {
    float f = (float)((d + 0.5) * scale - 0.5);
    int s = (int)floorf(f);
    f -= (float)s;
    if (horizontal) {                      // the fraction is reset at the borders on the x axis only
      if (s < 0) { s = 0; f = 0.f; }
      if (s >= src - 1) { s = src - 1; f = 0.f; }
      lo[d] = s; hi[d] = s + 1 < src ? s + 1 : src - 1;
    } else {
      lo[d] = s < 0 ? 0 : (s > src - 1 ? src - 1 : s);
      hi[d] = s + 1 < 0 ? 0 : (s + 1 > src - 1 ? src - 1 : s + 1);
    }
    k[2 * d] = (int)lrintf((1.f - f) * 2048.f);       // saturate_cast<short>(float): round half to even
    k[2 * d + 1] = (int)lrintf(f * 2048.f);
  }
}

// ---- Pillow bicubic (a = -0.5)
double bicubic(double x) {
  const double a = -0.5;
  if (x < 0.0) x = -x;
  if (x < 1.0) return ((a + 2.0) * x - (a + 3.0)) * x * x + 1;
  if (x < 2.0) return (((x - 5) * x + 8) * x - 4) * a;
  return 0.0;
}
int pil_axis(int src, int dst, int pad_to, std::vector<int>& lo, std::vector<int>& cnt, std::vector<int>& k) {
  double scale = (double)src / (double)dst, filterscale = scale;
  if (filterscale < 1.0) filterscale = 1.0;
  const double support = 2.0 * filterscale;
  const int taps = (int)ceil(support) * 2 + 1;
  const int ksize = (taps + pad_to - 1) / pad_to * pad_to;     // row length of the table, zero-filled past the taps
  lo.resize(dst); cnt.resize(dst); k.assign((size_t)dst * ksize, 0);
  const double ss = 1.0 / filterscale;
  std::vector<double> w(ksize);
  for (int xx = 0; xx < dst; ++xx) {
    const double center = (xx + 0.5) * scale;
    int xmin = (int)(center - support + 0.5);
    if (xmin < 0) xmin = 0;
    int xmax = (int)(center + support + 0.5);
    if (xmax > src) xmax = src;
    xmax -= xmin;
    double ww = 0.0;
    for (int x = 0; x < xmax; ++x) { w[x] = bicubic((x + xmin - center + 0.5) * ss); ww += w[x]; }
    for (int x = 0; x < xmax; ++x) {
      const double v = ww != 0.0 ? w[x] / ww : w[x];
      k[(size_t)xx * ksize + x] = v < 0 ? (int)(-0.5 + v * (1 << kPilBits)) : (int)(0.5 + v * (1 << kPilBits));
    }
    lo[xx] = xmin; cnt[xx] = xmax;
  }
  return ksize;
}

__device__ __forceinline__ uint32_t ingest_ldw(const uint32_t* p, uintptr_t end_up) {
  return reinterpret_cast<uintptr_t>(p) < end_up ? __ldg(p) : 0u;     // words past the source only sit under zero taps
}

// ---- flavour 0.  One thread = four consecutive canvas pixels of one row, all three colour planes.  The two taps of a row
// are neighbouring pixels = six consecutive bytes: read as aligned words and funnel-shifted (at the right border the
// second tap carries a zero coefficient, so what lies under it does not matter).
__global__ void __launch_bounds__(384) ingest_cv2_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ out,
                                                         IngestTables t) {
  const int X0 = threadIdx.x * 4, Y = blockIdx.x * blockDim.y + threadIdx.y, n = blockIdx.y;
  if (X0 >= t.out || Y >= t.out) return;
  const size_t frame_bytes = (size_t)t.h * t.w * 3;
  const uint8_t* img = src + (size_t)n * frame_bytes;
  const uintptr_t end_up = (reinterpret_cast<uintptr_t>(src + (size_t)gridDim.y * frame_bytes) + 3) & ~(uintptr_t)3;
  const int dy = Y - t.top;
  uint32_t px[3] = {0u, 0u, 0u};             // four packed bytes per output plane
  if (dy >= 0 && dy < t.new_h) {
    int r0, r1, b0, b1;
    if (t.area_fast) { r0 = 2 * dy; r1 = 2 * dy + 1; b0 = b1 = 0; }
    else { r0 = __ldg(t.y_lo + dy); r1 = __ldg(t.y_cnt + dy); b0 = __ldg(t.y_k + 2 * dy); b1 = __ldg(t.y_k + 2 * dy + 1); }
    const uint8_t* row0 = img + (size_t)r0 * t.w * 3;
    const uint8_t* row1 = img + (size_t)r1 * t.w * 3;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int dx = X0 + i - t.left;
      if (dx < 0 || dx >= t.new_w) continue;
      int c0, a0, a1;                        // the second tap is pixel c0 + 1 (or carries a zero coefficient)
      if (t.area_fast) { c0 = 2 * dx; a0 = a1 = 0; }
      else { c0 = __ldg(t.x_lo + dx); a0 = __ldg(t.x_k + 2 * dx); a1 = __ldg(t.x_k + 2 * dx + 1); }
      uint32_t t0[2], t1[2];                 // bytes 0..5 of the pixel pair in row0 / row1
      {
        const uintptr_t a0p = reinterpret_cast<uintptr_t>(row0 + c0 * 3), a1p = reinterpret_cast<uintptr_t>(row1 + c0 * 3);
        const uint32_t* w0 = reinterpret_cast<const uint32_t*>(a0p & ~(uintptr_t)3);
        const uint32_t* w1 = reinterpret_cast<const uint32_t*>(a1p & ~(uintptr_t)3);
        const int sh0 = (int)(a0p & 3) * 8, sh1 = (int)(a1p & 3) * 8;
        const uint32_t x0 = ingest_ldw(w0, end_up), x1 = ingest_ldw(w0 + 1, end_up), x2 = ingest_ldw(w0 + 2, end_up);
        const uint32_t y0 = ingest_ldw(w1, end_up), y1 = ingest_ldw(w1 + 1, end_up), y2 = ingest_ldw(w1 + 2, end_up);
        t0[0] = __funnelshift_r(x0, x1, sh0); t0[1] = __funnelshift_r(x1, x2, sh0);
        t1[0] = __funnelshift_r(y0, y1, sh1); t1[1] = __funnelshift_r(y1, y2, sh1);
      }
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const int s00 = (int)((t0[0] >> (8 * c)) & 0xff), s10 = (int)((t1[0] >> (8 * c)) & 0xff);
        const int s01 = c == 0 ? (int)(t0[0] >> 24) : (int)((t0[1] >> (8 * (c - 1))) & 0xff);
        const int s11 = c == 0 ? (int)(t1[0] >> 24) : (int)((t1[1] >> (8 * (c - 1))) & 0xff);
        int v;
        if (t.area_fast) v = (s00 + s01 + s10 + s11 + 2) >> 2;
        else {
          const int S0 = s00 * a0 + s01 * a1, S1 = s10 * a0 + s11 * a1;       // HResizeLinear, exact in int32
          v = (((b0 * (S0 >> 4)) >> 16) + ((b1 * (S1 >> 4)) >> 16) + 2) >> 2;  // VResizeLinear<uchar, int, short>
          v = v < 0 ? 0 : (v > 255 ? 255 : v);
        }
        px[2 - c] |= (uint32_t)v << (8 * i);     // BGR -> RGB: source channel c lands in plane 2 - c
      }
    }
  }
  const size_t plane = (size_t)t.out * t.out;
  uint8_t* o = out + (size_t)n * 3 * plane + (size_t)Y * t.out + X0;
#pragma unroll
  for (int c = 0; c < 3; ++c) *reinterpret_cast<uint32_t*>(o + c * plane) = px[c];
}

// ---- flavour 1.  One CTA = `th` consecutive canvas rows of one frame.  Horizontal pass of every source row under the
// tile's vertical taps into an 8-bit shared-memory image (Pillow rounds to 8 bits between the passes), then the
// vertical pass.  The kernel is bound by load instructions (L1 wavefronts), not by HBM, so both passes read 32-bit words:
// a horizontal work item = one output pixel (three channels) x four source rows; it streams the interleaved source
// bytes under its taps as aligned words (funnel-shifted to the pixel boundary), four pixels = three words at a time,
// with the four coefficients of the group in one 16-byte load.  The vertical pass owns four adjacent bytes of the
// horizontal image per thread.
__global__ void __launch_bounds__(384) ingest_pil_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ out,
                                                         IngestTables t) {
  extern __shared__ __align__(16) uint8_t hor[];     // [rows under the tile][hp], hp = new_w * 3 rounded up to 4
  constexpr int R = 4;
  const int Y0 = blockIdx.x * t.th, n = blockIdx.y;
  const size_t plane = (size_t)t.out * t.out;
  const int row_bytes = t.new_w * 3, hp = (row_bytes + 3) & ~3;
  int d0 = Y0 - t.top, d1 = Y0 + t.th - t.top;          // destination rows [d0, d1) of the resized image
  d0 = d0 < 0 ? 0 : d0; d1 = d1 > t.new_h ? t.new_h : d1;
  int ybase = 0;
  if (d0 < d1) {
    ybase = __ldg(t.y_lo + d0);
    const int nrows = __ldg(t.y_lo + d1 - 1) + __ldg(t.y_cnt + d1 - 1) - ybase;
    const size_t frame_bytes = (size_t)t.h * t.w * 3;
    const uint8_t* img = src + (size_t)n * frame_bytes;
    const uintptr_t end_up = (reinterpret_cast<uintptr_t>(src + (size_t)gridDim.y * frame_bytes) + 3) & ~(uintptr_t)3;
    const int items = t.new_w * ((nrows + R - 1) / R);
    for (int item = threadIdx.x; item < items; item += blockDim.x) {
      const int rg = item / t.new_w, x = item - rg * t.new_w, r0 = rg * R;
      const int lo = __ldg(t.x_lo + x), ng = (__ldg(t.x_cnt + x) + 3) >> 2;
      const int4* k4 = reinterpret_cast<const int4*>(t.x_k + (size_t)x * t.ksx);
      const uint32_t* wp[R];
      uint32_t carry[R], lo8[R][3], mid8[R][3];
      int sh[R], hi8[R][3];
#pragma unroll
      for (int u = 0; u < R; ++u) {
        const int row = r0 + u < nrows ? r0 + u : nrows - 1;         // rows past the tile: clamped, sums dropped
        const uintptr_t a = reinterpret_cast<uintptr_t>(img + ((size_t)(ybase + row) * t.w + lo) * 3);
        wp[u] = reinterpret_cast<const uint32_t*>(a & ~(uintptr_t)3);
        sh[u] = (int)(a & 3) * 8;
        carry[u] = ingest_ldw(wp[u], end_up);
#pragma unroll
        for (int c = 0; c < 3; ++c) { lo8[u][c] = mid8[u][c] = 0u; hi8[u][c] = 0; }
      }
      for (int g = 0; g < ng; ++g) {
        const int4 kk = __ldg(k4 + g);       // the four taps of the group as byte limbs: low, middle, high (signed)
#pragma unroll
        for (int u = 0; u < R; ++u) {
          const uint32_t l1 = ingest_ldw(wp[u] + 1, end_up), l2 = ingest_ldw(wp[u] + 2, end_up),
                         l3 = ingest_ldw(wp[u] + 3, end_up);
          wp[u] += 3;
          const uint32_t s0 = __funnelshift_r(carry[u], l1, sh[u]), s1 = __funnelshift_r(l1, l2, sh[u]),
                         s2 = __funnelshift_r(l2, l3, sh[u]);       // twelve bytes = four interleaved pixels
          carry[u] = l3;
          // the four bytes of one channel: stream bytes c, c + 3, c + 6, c + 9
          const uint32_t px[3] = {__byte_perm(__byte_perm(s0, s1, 0x0630), s2, 0x5210),
                                  __byte_perm(__byte_perm(s0, s1, 0x0741), s2, 0x6210),
                                  __byte_perm(__byte_perm(s0, s1, 0x0052), s2, 0x7410)};
#pragma unroll
          for (int c = 0; c < 3; ++c) {
            lo8[u][c] = __dp4a(px[c], (uint32_t)kk.x, lo8[u][c]);
            mid8[u][c] = __dp4a(px[c], (uint32_t)kk.y, mid8[u][c]);
            asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(hi8[u][c]) : "r"(px[c]), "r"(kk.z), "r"(hi8[u][c]));
          }
        }
      }
#pragma unroll
      for (int u = 0; u < R; ++u) {
        if (r0 + u < nrows) {
#pragma unroll
          for (int c = 0; c < 3; ++c) {
            // sum of byte * (low + 256 middle + 65536 high); exact modulo 2^32 and the true sum fits
            const int v = (int)(lo8[u][c] + (mid8[u][c] << 8) + ((uint32_t)hi8[u][c] << 16) + (1u << (kPilBits - 1))) >> kPilBits;
            hor[(r0 + u) * hp + x * 3 + c] = (uint8_t)(v < 0 ? 0 : (v > 255 ? 255 : v));
          }
        }
      }
    }
  }
  __syncthreads();
  const uint32_t* hor32 = reinterpret_cast<const uint32_t*>(hor);
  const int hp4 = hp >> 2, margin = t.out - t.new_w;
  for (int Y = Y0; Y < Y0 + t.th && Y < t.out; ++Y) {
    const int dy = Y - t.top;
    uint8_t* orow = out + (size_t)n * 3 * plane + (size_t)Y * t.out;
    if (dy < 0 || dy >= t.new_h) {                        // border row: zero words
      for (int j = threadIdx.x; j < 3 * (t.out >> 2); j += blockDim.x) {
        const int c = j / (t.out >> 2), q = j - c * (t.out >> 2);
        reinterpret_cast<uint32_t*>(orow + c * plane)[q] = 0u;
      }
      continue;
    }
    for (int j = threadIdx.x; j < 3 * margin; j += blockDim.x) {       // left / right border
      const int c = j / margin, m = j - c * margin;
      orow[c * plane + (m < t.left ? m : t.new_w + m)] = 0;
    }
    const int roff = __ldg(t.y_lo + dy) - ybase, ycnt = __ldg(t.y_cnt + dy);
    const int* k = t.y_k + (size_t)dy * t.ksy;
    for (int q = threadIdx.x; q < hp4; q += blockDim.x) {
      int a0 = 1 << (kPilBits - 1), a1 = a0, a2 = a0, a3 = a0;
      const uint32_t* hq = hor32 + roff * hp4 + q;
      for (int r = 0; r < ycnt; ++r) {
        const uint32_t v = hq[r * hp4];
        const int kv = __ldg(k + r);
        a0 += (int)(v & 0xff) * kv; a1 += (int)((v >> 8) & 0xff) * kv;
        a2 += (int)((v >> 16) & 0xff) * kv; a3 += (int)(v >> 24) * kv;
      }
      const int av[4] = {a0 >> kPilBits, a1 >> kPilBits, a2 >> kPilBits, a3 >> kPilBits};
#pragma unroll
      for (int b = 0; b < 4; ++b) {
        const int j = 4 * q + b;
        if (j < row_bytes) {
          const int x = j / 3, c = j - 3 * x;
          orow[c * plane + t.left + x] = (uint8_t)(av[b] < 0 ? 0 : (av[b] > 255 ? 255 : av[b]));
        }
      }
    }
  }
}

int build_tables(int h, int w, int flavour, int out, cudaStream_t st) {
  int device = 0;
  AHA_CUDA(cudaGetDevice(&device));
  for (const IngestTables& e : g_tabs)
    if (e.h == h && e.w == w && e.flavour == flavour && e.out == out && e.device == device) { g_tab = e; return AHA_OK; }
  if (g_tabs.size() >= kTableCacheEntries) {
    AHA_CUDA(cudaSetDevice(g_tabs.front().device));      // the oldest entry may belong to another device of the process
    AHA_CUDA(cudaDeviceSynchronize());
    cudaFree(g_tabs.front().block);
    AHA_CUDA(cudaSetDevice(device));
    g_tabs.erase(g_tabs.begin());
  }
  IngestTables t;
  t.h = h; t.w = w; t.flavour = flavour; t.out = out; t.device = device;
  // test/inference.py:540-545: the longer side becomes `out`, the other int((short / long) * out), in float64
  if (w > h) { t.new_w = out; t.new_h = (int)(((double)h / (double)w) * out); }
  else { t.new_h = out; t.new_w = (int)(((double)w / (double)h) * out); }
  AHA_REQUIRE(t.new_w >= 1 && t.new_h >= 1, "ingest_frames: the aspect ratio leaves an empty image");
  t.left = (out - t.new_w) / 2; t.top = (out - t.new_h) / 2;
  std::vector<int> xl, xc, xk, yl, yc, yk;
  if (flavour == 0) {
    t.area_fast = (w == 2 * t.new_w && h == 2 * t.new_h) ? 1 : 0;   // resize.cpp: INTER_LINEAR at exactly 2 x 2 -> INTER_AREA
    cv2_axis(w, t.new_w, true, xl, xc, xk);
    cv2_axis(h, t.new_h, false, yl, yc, yk);
    t.ksx = t.ksy = 2;
  } else {
    t.ksx = pil_axis(w, t.new_w, 4, xl, xc, xk);       // rows of four-tap groups, 16-byte aligned
    // the horizontal pass multiplies with dp4a: every group of four taps is stored as three words of byte limbs,
    // k = low + 256 * middle + 65536 * high with low, middle in [0, 255] and high a signed byte
    for (size_t g = 0; g + 3 < xk.size(); g += 4) {
      uint32_t lo8 = 0, mid8 = 0, hi8 = 0;
      for (int i = 0; i < 4; ++i) {
        const int k = xk[g + i], hi = k >> 16;
        AHA_REQUIRE(hi >= -128 && hi <= 127, "ingest_frames: coefficient out of the 24-bit range");
        lo8 |= (uint32_t)(k & 255) << (8 * i);
        mid8 |= (uint32_t)((k >> 8) & 255) << (8 * i);
        hi8 |= (uint32_t)(hi & 255) << (8 * i);
      }
      xk[g] = (int)lo8; xk[g + 1] = (int)mid8; xk[g + 2] = (int)hi8; xk[g + 3] = 0;
    }
    t.ksy = pil_axis(h, t.new_h, 1, yl, yc, yk);
    // rows per CTA: the largest tile whose horizontal-pass image stays under 100 KB (two CTAs per SM)
    for (int th = 16; th >= 1; th /= 2) {
      int worst = 0;
      for (int Y0 = 0; Y0 < out; Y0 += th) {
        int d0 = Y0 - t.top, d1 = Y0 + th - t.top;
        d0 = d0 < 0 ? 0 : d0; d1 = d1 > t.new_h ? t.new_h : d1;
        if (d0 < d1) worst = std::max(worst, yl[d1 - 1] + yc[d1 - 1] - yl[d0]);
      }
      t.th = th; t.tile_rows = worst;
      if ((size_t)worst * ((t.new_w * 3 + 3) & ~3) <= 100 * 1024) break;
    }
  }
  std::vector<int> host;
  auto put = [&](const std::vector<int>& v) {   // every table starts on a 16-byte boundary
    size_t o = host.size();
    host.insert(host.end(), v.begin(), v.end());
    host.resize((host.size() + 3) & ~(size_t)3, 0);
    return o;
  };
  const size_t o_xl = put(xl), o_xc = put(xc), o_xk = put(xk), o_yl = put(yl), o_yc = put(yc), o_yk = put(yk);
  AHA_CUDA(cudaMalloc(&t.block, host.size() * sizeof(int)));
  AHA_CUDA(cudaMemcpyAsync(t.block, host.data(), host.size() * sizeof(int), cudaMemcpyHostToDevice, st));
  AHA_CUDA(cudaStreamSynchronize(st));     // `host` is pageable memory that dies with this frame
  int* base = (int*)t.block;
  t.x_lo = base + o_xl; t.x_cnt = base + o_xc; t.x_k = base + o_xk;
  t.y_lo = base + o_yl; t.y_cnt = base + o_yc; t.y_k = base + o_yk;
  g_tab = t;
  g_tabs.push_back(t);
  return AHA_OK;
}

}  // namespace

int op_ingest_frames(const uint8_t* src_hwc, int n, int h, int w, int flavour, uint8_t* out_chw, int out, cudaStream_t st) {
  AHA_REQUIRE(flavour == 0 || flavour == 1, "ingest_frames: flavour must be 0 (cv2 bilinear, BGR) or 1 (PIL bicubic, RGB)");
  AHA_REQUIRE(n >= 0 && h >= 1 && w >= 1 && h <= 16384 && w <= 16384, "ingest_frames: bad frame geometry");
  AHA_REQUIRE(out >= 4 && out % 4 == 0 && out <= 1536, "ingest_frames: the canvas side must be a multiple of 4, at most 1536");
  AHA_REQUIRE((reinterpret_cast<uintptr_t>(out_chw) & 3) == 0, "ingest_frames: the canvas must be 4-byte aligned");
  // Pillow >= 11 resamples pictures more than 100 times taller than wide vertical pass first (Image.resize), the
  // reference's pinned 10.4.0 predates that; the result depends on the installed version, so the case is refused
  AHA_REQUIRE(!(flavour == 1 && (long long)h > 100LL * w && out < h),
              "ingest_frames: PIL flavour: height > 100 x width is resampled in a Pillow-version-dependent pass order; not supported");
  if (n == 0) return AHA_OK;
  AHA_REQUIRE(n <= 65535, "ingest_frames: at most 65535 frames per call");
  std::lock_guard<std::mutex> lock(g_tab_mutex);
  int rc = build_tables(h, w, flavour, out, st);
  if (rc != AHA_OK) return rc;
  const IngestTables t = g_tab;
  if (flavour == 0) {
    const int tx = out / 4, ty = tx <= 384 ? (384 / tx > 0 ? 384 / tx : 1) : 1;
    ingest_cv2_kernel<<<dim3((out + ty - 1) / ty, n), dim3(tx, ty), 0, st>>>(src_hwc, out_chw, t);
  } else {
    const size_t smem = (size_t)t.tile_rows * ((t.new_w * 3 + 3) & ~3);
    AHA_REQUIRE(smem <= 200 * 1024, "ingest_frames: the down-scale factor needs more than 200 KB of shared memory per row");
    if (smem > 48 * 1024)
      AHA_CUDA(cudaFuncSetAttribute(ingest_pil_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    ingest_pil_kernel<<<dim3((out + t.th - 1) / t.th, n), 384, smem, st>>>(src_hwc, out_chw, t);
  }
  ++g_aha_launch_count;
  AHA_CHECK_LAUNCH();
  return AHA_OK;
}
