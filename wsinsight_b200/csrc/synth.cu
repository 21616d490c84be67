// Synthetic in-memory slides (bench / test inputs) and the box-average thumbnail.
//
// The reference reads real slides through tiffslide/openslide (wsinsight/wsi.py:75-105,
// wsinsight/patchlib/pipeline.py:102-105); decoders are out of scope, so slides here are
// procedural: a counter-based integer hash drives a multi-octave value-noise "tissue" field
// (blobs with holes and islands), H&E-like colours on a near-white background and per-pixel
// noise.  The generator is integer-only and __host__ __device__, so the host can regenerate any
// window bit-exactly (CPU baseline / oracle inputs) without copying the 32 GB slide back.
#include "common.cuh"
#include "../../include/wsi_b200.h"

namespace {

__host__ __device__ __forceinline__ uint32_t mix32(uint32_t x) {
  x ^= x >> 16;
  x *= 0x7feb352dU;
  x ^= x >> 15;
  x *= 0x846ca68bU;
  x ^= x >> 16;
  return x;
}
__host__ __device__ __forceinline__ uint32_t hash3(uint32_t x, uint32_t y, uint32_t s) {
  return mix32(x * 0x9E3779B1U ^ mix32(y * 0x85EBCA77U ^ mix32(s + 0x27D4EB2FU)));
}

// Q16 smoothstep-interpolated value noise with lattice cell 2^lg pixels; returns [0, 65535]
__host__ __device__ __forceinline__ int vnoise(long long x, long long y, int lg, uint32_t seed) {
  const uint32_t cx = (uint32_t)(x >> lg), cy = (uint32_t)(y >> lg);
  const long long mask = (1ll << lg) - 1;
  long long fx = x & mask, fy = y & mask;
  int tx = lg >= 16 ? (int)(fx >> (lg - 16)) : (int)(fx << (16 - lg));
  int ty = lg >= 16 ? (int)(fy >> (lg - 16)) : (int)(fy << (16 - lg));
  // smoothstep in Q16: t*t*(3-2t)
  long long sx = ((long long)tx * tx >> 16) * ((3ll << 16) - 2ll * tx) >> 16;
  long long sy = ((long long)ty * ty >> 16) * ((3ll << 16) - 2ll * ty) >> 16;
  const int a = (int)(hash3(cx, cy, seed) & 0xFFFF), b = (int)(hash3(cx + 1, cy, seed) & 0xFFFF);
  const int c = (int)(hash3(cx, cy + 1, seed) & 0xFFFF), d = (int)(hash3(cx + 1, cy + 1, seed) & 0xFFFF);
  const long long top = a + (((long long)(b - a) * sx) >> 16);
  const long long bot = c + (((long long)(d - c) * sx) >> 16);
  return (int)(top + (((bot - top) * sy) >> 16));
}

__host__ __device__ __forceinline__ int base_lg(long long W, long long H) {
  long long m = W < H ? W : H;
  int lg = 0;
  while ((1ll << (lg + 1)) <= m) ++lg;
  lg -= 3;
  return lg < 5 ? 5 : lg;
}

__host__ __device__ __forceinline__ uint32_t synth_pixel(long long x, long long y, int lg, uint32_t seed) {
  // tissue field: 3 octaves, weights 4:2:1, range [0, 7*65535]
  const int f = 4 * vnoise(x, y, lg, seed) + 2 * vnoise(x, y, lg - 1, seed + 101) + vnoise(x, y, lg - 2, seed + 202);
  const uint32_t h = hash3((uint32_t)x, (uint32_t)y, seed ^ 0xA5A5A5A5U);
  int r, g, b;
  if (f > 7 * 32768 + 9000) {
    // H&E-like: pink/purple, modulated by a fine stain-density octave + pixel noise
    const int dens = vnoise(x, y, lg > 8 ? lg - 6 : 2, seed + 303) >> 8;   // 0..255
    r = 150 + (dens * 70 >> 8) + (int)(h & 31) - 16;
    g = 60 + (dens * 90 >> 8) + (int)((h >> 5) & 31) - 16;
    b = 140 + (dens * 60 >> 8) + (int)((h >> 10) & 31) - 16;
  } else {
    r = 243 + (int)(h & 3);
    g = 243 + (int)((h >> 2) & 3);
    b = 243 + (int)((h >> 4) & 3);
  }
  r = r < 0 ? 0 : (r > 255 ? 255 : r);
  g = g < 0 ? 0 : (g > 255 ? 255 : g);
  b = b < 0 ? 0 : (b > 255 ? 255 : b);
  return (uint32_t)r | ((uint32_t)g << 8) | ((uint32_t)b << 16) | 0xFF000000U;
}

__global__ void __launch_bounds__(256)
    k_synth(uint8_t* __restrict__ out, long long W, long long y_base, long long pitch, int lg, uint32_t seed) {
  const long long x4 = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  const long long y = y_base + blockIdx.y;
  if (x4 >= W) return;
  uint32_t* row = reinterpret_cast<uint32_t*>(out + y * pitch);
  if (x4 + 3 < W) {
    uint4 v;
    v.x = synth_pixel(x4, y, lg, seed);
    v.y = synth_pixel(x4 + 1, y, lg, seed);
    v.z = synth_pixel(x4 + 2, y, lg, seed);
    v.w = synth_pixel(x4 + 3, y, lg, seed);
    *reinterpret_cast<uint4*>(row + x4) = v;
  } else {
    for (long long x = x4; x < W; ++x) row[x] = synth_pixel(x, y, lg, seed);
  }
}

// one warp per thumbnail pixel; exact integer box mean with round-half-up
__global__ void __launch_bounds__(256)
    k_box_thumb(const uint8_t* __restrict__ src, long long W, long long H, long long pitch, uint8_t* __restrict__ dst,
                int Ht, int Wt) {
  const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= (long long)Ht * Wt) return;
  const int i = (int)(warp / Wt), j = (int)(warp - (long long)i * Wt);
  const long long x0 = j * W / Wt, x1 = (j + 1) * W / Wt;
  const long long y0 = i * H / Ht, y1 = (i + 1) * H / Ht;
  unsigned long long sr = 0, sg = 0, sbb = 0;
  for (long long y = y0; y < y1; ++y) {
    const uint32_t* row = reinterpret_cast<const uint32_t*>(src + y * pitch);
    for (long long x = x0 + lane; x < x1; x += 32) {
      const uint32_t p = __ldg(row + x);
      sr += p & 0xff;
      sg += (p >> 8) & 0xff;
      sbb += (p >> 16) & 0xff;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    sr += __shfl_xor_sync(0xffffffffu, sr, o);
    sg += __shfl_xor_sync(0xffffffffu, sg, o);
    sbb += __shfl_xor_sync(0xffffffffu, sbb, o);
  }
  if (lane == 0) {
    const unsigned long long cnt = (unsigned long long)((x1 - x0) * (y1 - y0));
    uint8_t* o = dst + ((size_t)i * Wt + j) * 3;
    if (cnt == 0) {
      o[0] = o[1] = o[2] = 0;
    } else {
      o[0] = (uint8_t)((sr + cnt / 2) / cnt);
      o[1] = (uint8_t)((sg + cnt / 2) / cnt);
      o[2] = (uint8_t)((sbb + cnt / 2) / cnt);
    }
  }
}

}  // namespace

extern "C" {

int wsi_synth_slide(uint8_t* d_rgba, long long W, long long H, long long pitch_bytes, unsigned seed, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  WSI_ARG_CHECK(d_rgba && W > 0 && H > 0 && pitch_bytes >= 4 * W && (pitch_bytes & 15) == 0 &&
                    ((uintptr_t)d_rgba & 15) == 0,
                "wsi_synth_slide: bad arguments");
  const int lg = base_lg(W, H);
  // grid.y is limited to 65535: launch in row bands (rows hash with their absolute y)
  const long long band = 65535;
  for (long long yb = 0; yb < H; yb += band) {
    const long long rows = H - yb < band ? H - yb : band;
    dim3 grid((unsigned)((W / 4 + 256) / 256), (unsigned)rows);
    k_synth<<<grid, 256, 0, stream>>>(d_rgba, W, yb, pitch_bytes, lg, seed);
    wsi_count_launch(1);
  }
  WSI_CUDA_CHECK(cudaGetLastError());
  return WSI_OK;
}

int wsi_synth_slide_host(uint8_t* h_rgba, long long x0, long long y0, int w, int h, long long W, long long H,
                         unsigned seed) {
  WSI_ARG_CHECK(h_rgba && w > 0 && h > 0 && W > 0 && H > 0, "wsi_synth_slide_host: bad arguments");
  const int lg = base_lg(W, H);
  uint32_t* o = reinterpret_cast<uint32_t*>(h_rgba);
  for (int yy = 0; yy < h; ++yy)
    for (int xx = 0; xx < w; ++xx) {
      const long long x = x0 + xx, y = y0 + yy;
      o[(size_t)yy * w + xx] = (x >= 0 && y >= 0 && x < W && y < H) ? synth_pixel(x, y, lg, seed) : 0u;
    }
  return WSI_OK;
}

int wsi_box_thumbnail(const uint8_t* d_rgba, long long W, long long H, long long pitch_bytes, uint8_t* d_rgb, int Ht,
                      int Wt, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  WSI_ARG_CHECK(d_rgba && d_rgb && W > 0 && H > 0 && Ht > 0 && Wt > 0 && Wt <= W && Ht <= H,
                "wsi_box_thumbnail: bad arguments");
  const long long warps = (long long)Ht * Wt;
  const long long blocks = (warps * 32 + 255) / 256;
  WSI_ARG_CHECK(blocks < (1ll << 31), "wsi_box_thumbnail: thumbnail too large");
  k_box_thumb<<<(unsigned)blocks, 256, 0, stream>>>(d_rgba, W, H, pitch_bytes, d_rgb, Ht, Wt);
  wsi_count_launch(1);
  WSI_CUDA_CHECK(cudaGetLastError());
  return WSI_OK;
}

}  // extern "C"
