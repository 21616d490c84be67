// N1: the thumbnail the reference segments, bit-exact with Pillow.
//
// Replaces (reference, huangch/wsinsight):
//   wsinsight/patchlib/pipeline.py:102-105   thumb = slide.get_thumbnail(thumbsize)
//     -> tiffslide 2.5.1 get_thumbnail: read level, composite on white, Image.thumbnail(size, LANCZOS)
//     -> Pillow Image.thumbnail / resize(reducing_gap=2.0): Reduce.c box reduction by an integer factor, then
//        Resample.c Lanczos-3 with the fractional source box, 22-bit fixed point, u8 intermediate.
//
// k_reduce   : the only stage that touches full-resolution pixels (32 GB on the C2 slide): HBM-bound, one
//              coalesced 16-byte load per thread and source row, column sums in registers, fx-wide sums through smem.
// k_resample_h / k_resample_v : Lanczos passes over the (small) reduced image.
#include "common.cuh"
#include "../../include/wsi_b200.h"
#include <math.h>
#include <vector>

namespace {

constexpr int PRECISION_BITS = 32 - 8 - 2;
constexpr int RED_THREADS = 256;

// Pillow Image.thumbnail(): aspect-preserving size
void thumb_size(long long w, long long h, int max_w, int max_h, int* ow, int* oh) {
  double x = (double)max_w, y = (double)max_h;
  if (x >= (double)w && y >= (double)h) {
    *ow = (int)w; *oh = (int)h;
    return;
  }
  const double aspect = (double)w / (double)h;
  auto pick = [](double number, auto key) {
    const double lo = floor(number), hi = ceil(number);
    double best = key(lo) <= key(hi) ? lo : hi;     // min() keeps the first (floor) on ties
    return best < 1.0 ? 1.0 : best;
  };
  if (x / y >= aspect) {
    x = pick(y * aspect, [&](double n) { return fabs(aspect - n / y); });
  } else {
    y = pick(x / aspect, [&](double n) { return n == 0.0 ? 0.0 : fabs(aspect - x / n); });
  }
  *ow = (int)x; *oh = (int)y;
}

uint32_t reduce_mult(int n) { return (uint32_t)((float)(1 << 24) / (float)n); }   // Reduce.c division_UINT32(n, 8)

double lanczos(double x) {
  auto sinc = [](double v) {
    if (v == 0.0) return 1.0;
    v = v * M_PI;
    return sin(v) / v;
  };
  if (-3.0 <= x && x < 3.0) return sinc(x) * sinc(x / 3.0);
  return 0.0;
}

struct Taps {
  std::vector<int> lo, cnt, coef;
  int ksize = 0;
};

// Resample.c precompute_coeffs (LANCZOS, support 3) + normalize_coeffs_8bpc; the box arrives as C floats
Taps lanczos_taps(int in_size, float in0f, float in1f, int out_size) {
  Taps t;
  const double in0 = (double)in0f;
  const double scale = (double)(in1f - in0f) / out_size;
  const double filterscale = scale < 1.0 ? 1.0 : scale;
  const double support = 3.0 * filterscale;
  const int ksize = (int)ceil(support) * 2 + 1;
  t.ksize = ksize;
  t.lo.resize(out_size);
  t.cnt.resize(out_size);
  t.coef.assign((size_t)out_size * ksize, 0);
  std::vector<double> k(ksize);
  const double ss = 1.0 / filterscale;
  for (int xx = 0; xx < out_size; ++xx) {
    const double center = in0 + (xx + 0.5) * scale;
    int xmin = (int)(center - support + 0.5);
    if (xmin < 0) xmin = 0;
    int xmax = (int)(center + support + 0.5);
    if (xmax > in_size) xmax = in_size;
    xmax -= xmin;
    double ww = 0.0;
    for (int x = 0; x < xmax; ++x) {
      k[x] = lanczos((x + xmin - center + 0.5) * ss);
      ww += k[x];
    }
    for (int x = 0; x < xmax; ++x) {
      if (ww != 0.0) k[x] /= ww;
      const double q = k[x] * (double)(1 << PRECISION_BITS);
      t.coef[(size_t)xx * ksize + x] = (int)(k[x] < 0 ? -0.5 + q : 0.5 + q);
    }
    t.lo[xx] = xmin;
    t.cnt[xx] = xmax;
  }
  return t;
}

struct RedArgs {
  const uint8_t* src;        // RGBA u8 [H][pitch]
  long long W, H, pitch;
  int fx, fy, Wr, Hr, outs;  // outs = reduced pixels per CTA (multiple of 4)
  uint32_t mult[2][2];       // [partial y][partial x] multipliers and half-counts
  uint32_t amend[2][2];
  uint32_t* dst;             // RGBX u8 [Hr][Wr]
};

// grid (ceil(Wr / outs), Hr); one CTA = `outs` reduced pixels of one reduced row
__global__ void __launch_bounds__(RED_THREADS) k_reduce(const RedArgs a) {
  extern __shared__ int s_col[];                   // [3][outs * fx] column sums
  const int seg = a.outs * a.fx;                   // source pixels per CTA row segment (multiple of 4)
  const long long x0 = (long long)blockIdx.x * seg;
  const long long y0 = (long long)blockIdx.y * a.fy;
  const int rows = (int)min((long long)a.fy, a.H - y0);
  for (int c4 = threadIdx.x; c4 * 4 < seg; c4 += RED_THREADS) {
    const long long x = x0 + 4 * c4;
    int s[12];
#pragma unroll
    for (int i = 0; i < 12; ++i) s[i] = 0;
    if (x < a.W) {
      const uint8_t* p = a.src + y0 * a.pitch + x * 4;
#pragma unroll 4
      for (int r = 0; r < rows; ++r) {
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(p + (long long)r * a.pitch));
        const uint32_t px[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          s[i] += (int)(px[i] & 0xff);
          s[4 + i] += (int)((px[i] >> 8) & 0xff);
          s[8 + i] += (int)((px[i] >> 16) & 0xff);
        }
      }
      // pixels at or beyond the slide width contribute nothing (the pitch padding may hold anything)
#pragma unroll
      for (int i = 0; i < 4; ++i)
        if (x + i >= a.W) s[i] = s[4 + i] = s[8 + i] = 0;
    }
#pragma unroll
    for (int ch = 0; ch < 3; ++ch)
      *reinterpret_cast<int4*>(s_col + ch * seg + 4 * c4) = make_int4(s[4 * ch], s[4 * ch + 1], s[4 * ch + 2], s[4 * ch + 3]);
  }
  __syncthreads();
  const int py = rows < a.fy ? 1 : 0;
  for (int o = threadIdx.x; o < a.outs; o += RED_THREADS) {
    const long long xo = (long long)blockIdx.x * a.outs + o;
    if (xo >= a.Wr) continue;
    const int cols = (int)min((long long)a.fx, a.W - xo * a.fx);
    const int px = cols < a.fx ? 1 : 0;
    uint32_t out = 0;
#pragma unroll
    for (int ch = 0; ch < 3; ++ch) {
      uint32_t ss = a.amend[py][px];
      const int* c = s_col + ch * seg + o * a.fx;
      for (int i = 0; i < cols; ++i) ss += (uint32_t)c[i];
      out |= ((ss * a.mult[py][px]) >> 24) << (8 * ch);
    }
    a.dst[(size_t)blockIdx.y * a.Wr + xo] = out;
  }
}

__device__ __forceinline__ uint32_t clip8(int v) {
  v >>= PRECISION_BITS;
  return (uint32_t)min(max(v, 0), 255);
}

// horizontal pass: src RGBX [Hs][Ws(pitch in pixels)] -> tmp RGBX [Hs][out_w]
__global__ void k_resample_h(const uint32_t* __restrict__ src, long long src_pitch_px, int Hs, int out_w,
                             const int* __restrict__ lo, const int* __restrict__ cnt, const int* __restrict__ coef,
                             int ksize, uint32_t* __restrict__ tmp) {
  const int xx = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (xx >= out_w || y >= Hs) return;
  const uint32_t* row = src + (size_t)y * src_pitch_px + lo[xx];
  const int* k = coef + (size_t)xx * ksize;
  const int n = cnt[xx];
  int s0 = 1 << (PRECISION_BITS - 1), s1 = s0, s2 = s0;
  for (int t = 0; t < n; ++t) {
    const uint32_t p = row[t];
    const int c = k[t];
    s0 += (int)(p & 0xff) * c;
    s1 += (int)((p >> 8) & 0xff) * c;
    s2 += (int)((p >> 16) & 0xff) * c;
  }
  tmp[(size_t)y * out_w + xx] = clip8(s0) | (clip8(s1) << 8) | (clip8(s2) << 16);
}

// vertical pass: tmp RGBX [Hs][out_w] -> out RGB u8 [out_h][out_w][3]
__global__ void k_resample_v(const uint32_t* __restrict__ tmp, int out_w, int out_h, const int* __restrict__ lo,
                             const int* __restrict__ cnt, const int* __restrict__ coef, int ksize,
                             uint8_t* __restrict__ out) {
  const int xx = blockIdx.x * blockDim.x + threadIdx.x, yy = blockIdx.y;
  if (xx >= out_w || yy >= out_h) return;
  const uint32_t* col = tmp + (size_t)lo[yy] * out_w + xx;
  const int* k = coef + (size_t)yy * ksize;
  const int n = cnt[yy];
  int s0 = 1 << (PRECISION_BITS - 1), s1 = s0, s2 = s0;
  for (int t = 0; t < n; ++t) {
    const uint32_t p = col[(size_t)t * out_w];
    const int c = k[t];
    s0 += (int)(p & 0xff) * c;
    s1 += (int)((p >> 8) & 0xff) * c;
    s2 += (int)((p >> 16) & 0xff) * c;
  }
  uint8_t* o = out + ((size_t)yy * out_w + xx) * 3;
  o[0] = (uint8_t)clip8(s0);
  o[1] = (uint8_t)clip8(s1);
  o[2] = (uint8_t)clip8(s2);
}

// RGBA [H][pitch] -> packed RGB (no resampling needed: the slide already fits the requested size)
__global__ void k_rgba_to_rgb(const uint8_t* __restrict__ src, long long pitch, int W, int H, uint8_t* __restrict__ out) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= W || y >= H) return;
  const uint8_t* p = src + (size_t)y * pitch + (size_t)x * 4;
  uint8_t* o = out + ((size_t)y * W + x) * 3;
  o[0] = p[0]; o[1] = p[1]; o[2] = p[2];
}

}  // namespace

extern "C" {

// Pillow's thumbnail geometry for a W x H slide and a requested (max_w, max_h): output size and reduce factors.
int wsi_thumbnail_geometry(long long W, long long H, int max_w, int max_h, int* out_w, int* out_h, int* fx, int* fy) {
  WSI_ARG_CHECK(W >= 1 && H >= 1 && max_w >= 1 && max_h >= 1 && out_w && out_h, "wsi_thumbnail_geometry: bad arguments");
  int ow, oh;
  thumb_size(W, H, max_w, max_h, &ow, &oh);
  *out_w = ow; *out_h = oh;
  int rx = (int)((double)W / ow / 2.0), ry = (int)((double)H / oh / 2.0);   // Image.resize(reducing_gap=2.0)
  if (rx < 1) rx = 1;
  if (ry < 1) ry = 1;
  if (fx) *fx = rx;
  if (fy) *fy = ry;
  return WSI_OK;
}

// device scratch: reduced image + horizontal-pass intermediate (both RGBX u32) + the two tap tables
size_t wsi_thumbnail_workspace_bytes(long long W, long long H, int max_w, int max_h) {
  int ow, oh, fx, fy;
  if (wsi_thumbnail_geometry(W, H, max_w, max_h, &ow, &oh, &fx, &fy)) return 0;
  const long long Wr = (W + fx - 1) / fx, Hr = (H + fy - 1) / fy;
  const bool reduced = fx > 1 || fy > 1;
  size_t b = 0;
  if (reduced) b += (size_t)Wr * Hr * 4;
  b += (size_t)Hr * ow * 4;
  const double sx = (double)W / fx / ow, sy = (double)H / fy / oh;
  const int kx = (int)ceil(3.0 * (sx < 1 ? 1 : sx)) * 2 + 1 + 2, ky = (int)ceil(3.0 * (sy < 1 ? 1 : sy)) * 2 + 1 + 2;
  b += ((size_t)ow * (2 + kx) + (size_t)oh * (2 + ky)) * 4 + 1024;
  return b;
}

// RGBA u8 slide (alpha assumed opaque: compositing on the white background is the identity) -> RGB u8
// [out_h][out_w][3], bit-exact with Image.thumbnail((max_w, max_h), LANCZOS) of the RGB image.
int wsi_lanczos_thumbnail(const uint8_t* d_slide_rgba, long long W, long long H, long long pitch_bytes, int max_w,
                          int max_h, uint8_t* d_out_rgb, void* d_workspace, size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  WSI_ARG_CHECK(d_slide_rgba && d_out_rgb, "wsi_lanczos_thumbnail: null pointer");
  WSI_ARG_CHECK(((uintptr_t)d_slide_rgba & 15) == 0 && (pitch_bytes & 15) == 0 && pitch_bytes >= W * 4,
                "wsi_lanczos_thumbnail: slide base/pitch must be 16-byte aligned RGBA");
  int ow, oh, fx, fy;
  int rc = wsi_thumbnail_geometry(W, H, max_w, max_h, &ow, &oh, &fx, &fy);
  if (rc) return rc;
  if ((long long)ow == W && (long long)oh == H) {
    dim3 grid((unsigned)((W + 255) / 256), (unsigned)H);
    k_rgba_to_rgb<<<grid, 256, 0, stream>>>(d_slide_rgba, pitch_bytes, (int)W, (int)H, d_out_rgb);
    wsi_count_launch(1);
    WSI_CUDA_CHECK(cudaGetLastError());
    return WSI_OK;
  }
  // Pillow reduces 1xN / Nx1 with dedicated loops whose rounding this kernel does not restate
  WSI_ARG_CHECK((fx == 1) == (fy == 1), "wsi_lanczos_thumbnail: reduce factor (%d, %d) with a single unit axis is unsupported",
                fx, fy);
  WSI_ARG_CHECK(d_workspace && workspace_bytes >= wsi_thumbnail_workspace_bytes(W, H, max_w, max_h),
                "wsi_lanczos_thumbnail: workspace too small");
  const bool reduced = fx > 1;
  const long long Wr = (W + fx - 1) / fx, Hr = (H + fy - 1) / fy;
  uint8_t* ws = (uint8_t*)d_workspace;
  uint32_t* d_red = nullptr;
  if (reduced) {
    d_red = (uint32_t*)ws;
    ws += (size_t)Wr * Hr * 4;
  }
  uint32_t* d_tmp = (uint32_t*)ws;
  ws += (size_t)Hr * ow * 4;
  int launches = 0;
  if (reduced) {
    RedArgs a;
    a.src = d_slide_rgba; a.W = W; a.H = H; a.pitch = pitch_bytes;
    a.fx = fx; a.fy = fy; a.Wr = (int)Wr; a.Hr = (int)Hr;
    int outs = 1024 / fx / 4 * 4;
    if (outs < 4) outs = 4;
    if (outs > 64) outs = 64;
    a.outs = outs;
    const int pxn = (int)(W % fx), pyn = (int)(H % fy);
    for (int py = 0; py < 2; ++py)
      for (int px = 0; px < 2; ++px) {
        const int n = (py ? (pyn ? pyn : fy) : fy) * (px ? (pxn ? pxn : fx) : fx);
        a.mult[py][px] = reduce_mult(n);
        a.amend[py][px] = (uint32_t)(n / 2);
      }
    a.dst = d_red;
    const size_t smem = (size_t)3 * outs * fx * sizeof(int);
    WSI_ARG_CHECK(smem <= 160 * 1024, "wsi_lanczos_thumbnail: reduce factor %d too large", fx);
    if (smem > 48 * 1024)
      WSI_CUDA_CHECK(cudaFuncSetAttribute(k_reduce, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    dim3 grid((unsigned)((Wr + outs - 1) / outs), (unsigned)Hr);
    k_reduce<<<grid, RED_THREADS, smem, stream>>>(a);
    ++launches;
  }
  // Lanczos passes over the reduced image (or the slide itself when no reduction applies)
  const int Ws = (int)Wr, Hs = (int)Hr;
  Taps tx = lanczos_taps(Ws, 0.0f, (float)((double)W / fx), ow);
  Taps ty = lanczos_taps(Hs, 0.0f, (float)((double)H / fy), oh);
  int* d_tab = (int*)(((uintptr_t)ws + 255) & ~(uintptr_t)255);
  int* d_xlo = d_tab;
  int* d_xcnt = d_xlo + ow;
  int* d_xk = d_xcnt + ow;
  int* d_ylo = d_xk + (size_t)ow * tx.ksize;
  int* d_ycnt = d_ylo + oh;
  int* d_yk = d_ycnt + oh;
  WSI_ARG_CHECK((uint8_t*)(d_yk + (size_t)oh * ty.ksize) <= (uint8_t*)d_workspace + workspace_bytes,
                "wsi_lanczos_thumbnail: workspace too small for the tap tables");
  // pageable-host copies are staged by the runtime before cudaMemcpyAsync returns: the vectors may go out of scope
  WSI_CUDA_CHECK(cudaMemcpyAsync(d_xlo, tx.lo.data(), sizeof(int) * ow, cudaMemcpyHostToDevice, stream));
  WSI_CUDA_CHECK(cudaMemcpyAsync(d_xcnt, tx.cnt.data(), sizeof(int) * ow, cudaMemcpyHostToDevice, stream));
  WSI_CUDA_CHECK(cudaMemcpyAsync(d_xk, tx.coef.data(), sizeof(int) * tx.coef.size(), cudaMemcpyHostToDevice, stream));
  WSI_CUDA_CHECK(cudaMemcpyAsync(d_ylo, ty.lo.data(), sizeof(int) * oh, cudaMemcpyHostToDevice, stream));
  WSI_CUDA_CHECK(cudaMemcpyAsync(d_ycnt, ty.cnt.data(), sizeof(int) * oh, cudaMemcpyHostToDevice, stream));
  WSI_CUDA_CHECK(cudaMemcpyAsync(d_yk, ty.coef.data(), sizeof(int) * ty.coef.size(), cudaMemcpyHostToDevice, stream));
  const uint32_t* src = reduced ? d_red : reinterpret_cast<const uint32_t*>(d_slide_rgba);
  const long long src_pitch_px = reduced ? Wr : pitch_bytes / 4;
  {
    dim3 grid((unsigned)((ow + 127) / 128), (unsigned)Hs);
    k_resample_h<<<grid, 128, 0, stream>>>(src, src_pitch_px, Hs, ow, d_xlo, d_xcnt, d_xk, tx.ksize, d_tmp);
    ++launches;
  }
  {
    dim3 grid((unsigned)((ow + 127) / 128), (unsigned)oh);
    k_resample_v<<<grid, 128, 0, stream>>>(d_tmp, ow, oh, d_ylo, d_ycnt, d_yk, ty.ksize, d_out_rgb);
    ++launches;
  }
  wsi_count_launch(launches);
  WSI_CUDA_CHECK(cudaGetLastError());
  return WSI_OK;
}

}  // extern "C"
