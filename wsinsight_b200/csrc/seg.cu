// K1: tissue segmentation of slide thumbnails (batched over a cohort), K2: in-tissue patch-grid filter.
//
// Replaces (reference, huangch/wsinsight):
//   wsinsight/patchlib/segment.py:54-97   cv.cvtColor/medianBlur/threshold, skimage
//                                         binary_closing/remove_small_objects/remove_small_holes
//   wsinsight/patchlib/patch.py:57-130    cv.findContours + Shapely union/difference
//   wsinsight/patchlib/patch.py:210-242   itertools grid + STRtree.query(contains) + sort
//
// All arithmetic is integer and bit-exact with the CPU oracle (oracle/segment.py,
// oracle/patchgrid.py).  HBM/latency-bound byte work: no tensor cores here.  Every stage is one
// launch over ALL thumbnails of the cohort (flat tile / pixel index -> slide by binary search over
// a prefix table), so a 64-slide cohort costs the same number of launches as one slide.
#include "common.cuh"
#include "../../include/wsi_b200.h"
#include <stdlib.h>
#include <mutex>
#include <vector>

namespace {

struct SegSlide {
  int H, W, pitch;
  int min_obj, min_hole;
  int tiles_x;            // tiles per row for the 64x16 tiling (K1a/K1b)
  int ctiles_x;           // tiles per row for the 32x16 tiling (CCL)
  int pad_;
  long long rgb_off;      // bytes from d_rgb
  long long mask_off;     // bytes from d_mask / d_outer / tmp
  long long lab_off;      // ints from L / aux: slide s owns n_s + 1 entries (last = frame node)
};

struct SegBatch {
  const SegSlide* slides;      // device
  const long long* tile_base;  // [n+1] prefix of 64x16 tiles
  const long long* ctile_base; // [n+1] prefix of 32x16 tiles
  const long long* px_base;    // [n+1] prefix of pixels
  int n;
};

__device__ __forceinline__ int find_slide(const long long* __restrict__ base, int n, long long idx) {
  int lo = 0, hi = n - 1;
  while (lo < hi) {
    int mid = (lo + hi + 1) >> 1;
    if (base[mid] <= idx) lo = mid; else hi = mid - 1;
  }
  return lo;
}

// Per-pixel kernels: the slide of a CTA's first pixel is searched once (thread 0) and broadcast; a pixel past the end
// of that slide steps forward linearly (a CTA straddles at most a couple of slides).  The binary search per thread
// was a chain of log2(n) dependent loads in front of every pixel of a cohort.  Must be called by ALL threads of the CTA.
__device__ __forceinline__ int find_slide_cta(const long long* __restrict__ base, int n, long long g) {
  __shared__ int s_first;
  if (threadIdx.x == 0) s_first = find_slide(base, n, (long long)blockIdx.x * blockDim.x);
  __syncthreads();
  int s = s_first;
  while (s + 1 < n && base[s + 1] <= g) ++s;
  return s;
}

// ----------------------------------------------------------------------------
// K1a: RGB -> HSV saturation (OpenCV 8-bit formula) -> (S > thresh) -> k x k box
// count with replicate border >= (k*k+1)/2   [== medianBlur(S,k) > thresh]
// ----------------------------------------------------------------------------
__constant__ int c_sdiv[256];

constexpr int K1_TW = 64;
constexpr int K1_TH = 16;
constexpr int K1_THREADS = 256;

__device__ __forceinline__ int sat_of(const uint8_t* __restrict__ p) {
  const int R = p[0], G = p[1], B = p[2];
  const int v = max(R, max(G, B));
  const int d = v - min(R, min(G, B));
  return (d * c_sdiv[v] + (1 << 11)) >> 12;
}

// legacy form (k*k taps per output), kept for A/B (WSI_B200_SEG_LEGACY=1)
__global__ void __launch_bounds__(K1_THREADS)
    k_sat_median_thresh_v1(const uint8_t* __restrict__ rgb_all, SegBatch sb, int k, int thresh,
                           uint8_t* __restrict__ out_all) {
  extern __shared__ uint8_t s_b[];  // (TH+2r) x (TW+2r) thresholded bits
  const int s = find_slide(sb.tile_base, sb.n, blockIdx.x);
  const SegSlide sl = sb.slides[s];
  const int tile = (int)(blockIdx.x - sb.tile_base[s]);
  const int ty = tile / sl.tiles_x, tx = tile - ty * sl.tiles_x;
  const uint8_t* __restrict__ rgb = rgb_all + sl.rgb_off;
  uint8_t* __restrict__ out = out_all + sl.mask_off;
  const int H = sl.H, W = sl.W;
  const int r = k >> 1;
  const int sw = K1_TW + 2 * r, sh = K1_TH + 2 * r;
  const int x0 = tx * K1_TW - r, y0 = ty * K1_TH - r;
  for (int i = threadIdx.x; i < sw * sh; i += K1_THREADS) {
    int ly = i / sw, lx = i - ly * sw;
    int gx = min(max(x0 + lx, 0), W - 1);   // BORDER_REPLICATE
    int gy = min(max(y0 + ly, 0), H - 1);
    s_b[i] = (sat_of(rgb + (size_t)gy * sl.pitch + (size_t)gx * 3) > thresh) ? 1 : 0;
  }
  __syncthreads();
  const int need = (k * k + 1) >> 1;
  for (int i = threadIdx.x; i < K1_TW * K1_TH; i += K1_THREADS) {
    int ly = i / K1_TW, lx = i - ly * K1_TW;
    int gx = tx * K1_TW + lx, gy = ty * K1_TH + ly;
    if (gx >= W || gy >= H) continue;
    int cnt = 0;
    for (int dy = 0; dy < k; ++dy) {
      const uint8_t* row = s_b + (ly + dy) * sw + lx;
      for (int dx = 0; dx < k; ++dx) cnt += row[dx];
    }
    out[(size_t)gy * W + gx] = cnt >= need ? 1 : 0;
  }
}

// separable form: threshold bits -> k-wide row sums -> k-high column sums (2k taps per output)
__global__ void __launch_bounds__(K1_THREADS)
    k_sat_median_thresh(const uint8_t* __restrict__ rgb_all, SegBatch sb, int k, int thresh,
                        uint8_t* __restrict__ out_all) {
  extern __shared__ uint8_t s_b[];  // (TH+2r) x (TW+2r) thresholded bits, then (TH+2r) x TW row sums
  const int s = find_slide(sb.tile_base, sb.n, blockIdx.x);
  const SegSlide sl = sb.slides[s];
  const int tile = (int)(blockIdx.x - sb.tile_base[s]);
  const int ty = tile / sl.tiles_x, tx = tile - ty * sl.tiles_x;
  const uint8_t* __restrict__ rgb = rgb_all + sl.rgb_off;
  uint8_t* __restrict__ out = out_all + sl.mask_off;
  const int H = sl.H, W = sl.W;
  const int r = k >> 1;
  const int sw = K1_TW + 2 * r, sh = K1_TH + 2 * r;
  uint8_t* s_h = s_b + sw * sh;
  const int x0 = tx * K1_TW - r, y0 = ty * K1_TH - r;
  for (int i = threadIdx.x; i < sw * sh; i += K1_THREADS) {
    int ly = i / sw, lx = i - ly * sw;
    int gx = min(max(x0 + lx, 0), W - 1);   // BORDER_REPLICATE
    int gy = min(max(y0 + ly, 0), H - 1);
    s_b[i] = (sat_of(rgb + (size_t)gy * sl.pitch + (size_t)gx * 3) > thresh) ? 1 : 0;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < sh * K1_TW; i += K1_THREADS) {
    const int ly = i / K1_TW, lx = i - ly * K1_TW;
    const uint8_t* row = s_b + ly * sw + lx;
    int c = 0;
    for (int dx = 0; dx < k; ++dx) c += row[dx];
    s_h[i] = (uint8_t)c;                     // k <= 31
  }
  __syncthreads();
  const int need = (k * k + 1) >> 1;
  for (int i = threadIdx.x; i < K1_TW * K1_TH; i += K1_THREADS) {
    int ly = i / K1_TW, lx = i - ly * K1_TW;
    int gx = tx * K1_TW + lx, gy = ty * K1_TH + ly;
    if (gx >= W || gy >= H) continue;
    int cnt = 0;
    for (int dy = 0; dy < k; ++dy) cnt += s_h[(ly + dy) * K1_TW + lx];
    out[(size_t)gy * W + gx] = cnt >= need ? 1 : 0;
  }
}

// Otsu mode (extra, not upstream: SURVEY F2 / north_star "thumbnail Otsu threshold"): the exact k x k median of S
// (cv.medianBlur, BORDER_REPLICATE) by 8-step bisection on the value, plus a per-slide 256-bin histogram.
__global__ void __launch_bounds__(K1_THREADS)
    k_sat_median_value(const uint8_t* __restrict__ rgb_all, SegBatch sb, int k, uint8_t* __restrict__ val_all,
                       int* __restrict__ hist_all) {
  extern __shared__ uint8_t s_b[];  // (TH+2r) x (TW+2r) saturation values
  __shared__ int s_hist[256];
  const int s = find_slide(sb.tile_base, sb.n, blockIdx.x);
  const SegSlide sl = sb.slides[s];
  const int tile = (int)(blockIdx.x - sb.tile_base[s]);
  const int ty = tile / sl.tiles_x, tx = tile - ty * sl.tiles_x;
  const uint8_t* __restrict__ rgb = rgb_all + sl.rgb_off;
  uint8_t* __restrict__ out = val_all + sl.mask_off;
  const int H = sl.H, W = sl.W;
  const int r = k >> 1;
  const int sw = K1_TW + 2 * r, sh = K1_TH + 2 * r;
  const int x0 = tx * K1_TW - r, y0 = ty * K1_TH - r;
  s_hist[threadIdx.x] = 0;                  // K1_THREADS == 256
  for (int i = threadIdx.x; i < sw * sh; i += K1_THREADS) {
    int ly = i / sw, lx = i - ly * sw;
    int gx = min(max(x0 + lx, 0), W - 1);
    int gy = min(max(y0 + ly, 0), H - 1);
    s_b[i] = (uint8_t)sat_of(rgb + (size_t)gy * sl.pitch + (size_t)gx * 3);
  }
  __syncthreads();
  const int rank = (k * k) >> 1;            // 0-based index of the median in the sorted window
  for (int i = threadIdx.x; i < K1_TW * K1_TH; i += K1_THREADS) {
    int ly = i / K1_TW, lx = i - ly * K1_TW;
    int gx = tx * K1_TW + lx, gy = ty * K1_TH + ly;
    if (gx >= W || gy >= H) continue;
    int med = 0;                            // largest c with #{window < c} <= rank
    for (int bit = 7; bit >= 0; --bit) {
      const int c = med | (1 << bit);
      int cnt = 0;
      for (int dy = 0; dy < k; ++dy) {
        const uint8_t* row = s_b + (ly + dy) * sw + lx;
        for (int dx = 0; dx < k; ++dx) cnt += row[dx] < c;
      }
      if (cnt <= rank) med = c;
    }
    out[(size_t)gy * W + gx] = (uint8_t)med;
    atomicAdd(&s_hist[med], 1);
  }
  __syncthreads();
  const int hv = s_hist[threadIdx.x];
  if (hv) atomicAdd(&hist_all[s * 256 + threadIdx.x], hv);
}

// cv::threshold(..., THRESH_OTSU) threshold selection (OpenCV imgproc/thresh.cpp getThreshVal_Otsu_8u): the same
// sequential float64 operations in the same order, no FMA contraction.  One thread per slide.
__global__ void k_otsu(const int* __restrict__ hist_all, SegBatch sb, int* __restrict__ thr_out) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= sb.n) return;
  const int* __restrict__ h = hist_all + s * 256;
  const SegSlide sl = sb.slides[s];
  const double scale = __ddiv_rn(1.0, (double)(sl.W * sl.H));
  double mu = 0.0;
  for (int i = 0; i < 256; ++i) mu = __dadd_rn(mu, __dmul_rn((double)i, (double)h[i]));
  mu = __dmul_rn(mu, scale);
  const double eps = 1.1920928955078125e-07;   // FLT_EPSILON
  double mu1 = 0.0, q1 = 0.0, max_sigma = 0.0;
  int max_val = 0;
  for (int i = 0; i < 256; ++i) {
    const double p_i = __dmul_rn((double)h[i], scale);
    mu1 = __dmul_rn(mu1, q1);
    q1 = __dadd_rn(q1, p_i);
    const double q2 = __dsub_rn(1.0, q1);
    if (fmin(q1, q2) < eps || fmax(q1, q2) > __dsub_rn(1.0, eps)) continue;
    mu1 = __ddiv_rn(__dadd_rn(mu1, __dmul_rn((double)i, p_i)), q1);
    const double mu2 = __ddiv_rn(__dsub_rn(mu, __dmul_rn(q1, mu1)), q2);
    const double d = __dsub_rn(mu1, mu2);
    const double sigma = __dmul_rn(__dmul_rn(__dmul_rn(q1, q2), d), d);
    if (sigma > max_sigma) {
      max_sigma = sigma;
      max_val = i;
    }
  }
  thr_out[s] = max_val;
}

// mask = value > thr[slide]   (cv.threshold THRESH_BINARY with the per-slide Otsu threshold)
__global__ void __launch_bounds__(256)
    k_thresh_slide(const uint8_t* __restrict__ val_all, SegBatch sb, long long total, const int* __restrict__ thr,
                   uint8_t* __restrict__ out_all) {
  const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int s = find_slide_cta(sb.px_base, sb.n, g < total ? g : total - 1);
  if (g >= total) return;
  const SegSlide sl = sb.slides[s];
  const long long idx = g - sb.px_base[s];
  (out_all + sl.mask_off)[idx] = (val_all + sl.mask_off)[idx] > thr[s] ? 1 : 0;
}

// ----------------------------------------------------------------------------
// K1b: closing with ones(k,k):  dilation offsets [-da .. k-1-da] (outside = 0) then erosion offsets
// [-ea .. k-1-ea] (outside = 1).  (da, ea) is the footprint-origin convention; it only matters for even k
// and only within k/2 pixels of the image border (a closing is translation invariant in the interior):
//   scipy / raw even footprint (default, == end-padded):  da = (k-1)/2, ea = k/2
//   front-padded (k+1)x(k+1) footprint:                   da = k/2,     ea = (k-1)/2
// ----------------------------------------------------------------------------
// legacy form (k*k taps per stage), kept for A/B (WSI_B200_SEG_LEGACY=1)
__global__ void __launch_bounds__(K1_THREADS)
    k_closing_v1(const uint8_t* __restrict__ in_all, SegBatch sb, int k, int dlo, int elo,
                 uint8_t* __restrict__ out_all) {
  extern __shared__ uint8_t s_m[];
  const int s = find_slide(sb.tile_base, sb.n, blockIdx.x);
  const SegSlide sl = sb.slides[s];
  const int tile = (int)(blockIdx.x - sb.tile_base[s]);
  const int ty = tile / sl.tiles_x, tx = tile - ty * sl.tiles_x;
  const uint8_t* __restrict__ in = in_all + sl.mask_off;
  uint8_t* __restrict__ out = out_all + sl.mask_off;
  const int H = sl.H, W = sl.W;
  const int halo = k - 1;
  const int iw = K1_TW + 2 * halo, ih = K1_TH + 2 * halo;      // input tile
  const int dw = K1_TW + halo, dh = K1_TH + halo;              // dilated tile
  uint8_t* s_in = s_m;
  uint8_t* s_d = s_m + iw * ih;
  const int bx = tx * K1_TW, by = ty * K1_TH;
  // input tile origin: (bx - elo - dlo, by - elo - dlo)
  for (int i = threadIdx.x; i < iw * ih; i += K1_THREADS) {
    int ly = i / iw, lx = i - ly * iw;
    int gx = bx - elo - dlo + lx, gy = by - elo - dlo + ly;
    s_in[i] = (gx >= 0 && gx < W && gy >= 0 && gy < H) ? in[(size_t)gy * W + gx] : 0;
  }
  __syncthreads();
  // dilated tile origin: (bx - elo, by - elo); positions outside the image read as 1 for the erosion
  for (int i = threadIdx.x; i < dw * dh; i += K1_THREADS) {
    int ly = i / dw, lx = i - ly * dw;
    int gx = bx - elo + lx, gy = by - elo + ly;
    uint8_t v;
    if (gx < 0 || gx >= W || gy < 0 || gy >= H) {
      v = 1;
    } else {
      v = 0;
      for (int dy = 0; dy < k; ++dy) {
        const uint8_t* row = s_in + (ly + dy) * iw + lx;
        for (int dx = 0; dx < k; ++dx) v |= row[dx];
      }
    }
    s_d[i] = v;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < K1_TW * K1_TH; i += K1_THREADS) {
    int ly = i / K1_TW, lx = i - ly * K1_TW;
    int gx = bx + lx, gy = by + ly;
    if (gx >= W || gy >= H) continue;
    uint8_t v = 1;
    for (int dy = 0; dy < k; ++dy) {
      const uint8_t* row = s_d + (ly + dy) * dw + lx;
      for (int dx = 0; dx < k; ++dx) v &= row[dx];
    }
    out[(size_t)gy * W + gx] = v;
  }
}

// separable form: row OR -> column OR (+ outside = 1) -> row AND -> column AND  (4k taps per output)
__global__ void __launch_bounds__(K1_THREADS)
    k_closing(const uint8_t* __restrict__ in_all, SegBatch sb, int k, int dlo, int elo,
              uint8_t* __restrict__ out_all) {
  extern __shared__ uint8_t s_m[];
  const int s = find_slide(sb.tile_base, sb.n, blockIdx.x);
  const SegSlide sl = sb.slides[s];
  const int tile = (int)(blockIdx.x - sb.tile_base[s]);
  const int ty = tile / sl.tiles_x, tx = tile - ty * sl.tiles_x;
  const uint8_t* __restrict__ in = in_all + sl.mask_off;
  uint8_t* __restrict__ out = out_all + sl.mask_off;
  const int H = sl.H, W = sl.W;
  const int halo = k - 1;
  const int iw = K1_TW + 2 * halo, ih = K1_TH + 2 * halo;      // input tile
  const int dw = K1_TW + halo, dh = K1_TH + halo;              // dilated tile
  uint8_t* s_in = s_m;                 // ih x iw
  uint8_t* s_t = s_in + iw * ih;       // ih x dw   row OR
  uint8_t* s_d = s_t + ih * dw;        // dh x dw   dilated (outside the image = 1)
  uint8_t* s_e = s_d + dh * dw;        // dh x TW   row AND
  const int bx = tx * K1_TW, by = ty * K1_TH;
  for (int i = threadIdx.x; i < iw * ih; i += K1_THREADS) {
    int ly = i / iw, lx = i - ly * iw;
    int gx = bx - elo - dlo + lx, gy = by - elo - dlo + ly;
    s_in[i] = (gx >= 0 && gx < W && gy >= 0 && gy < H) ? in[(size_t)gy * W + gx] : 0;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < ih * dw; i += K1_THREADS) {
    int ly = i / dw, lx = i - ly * dw;
    const uint8_t* row = s_in + ly * iw + lx;
    uint8_t v = 0;
    for (int dx = 0; dx < k; ++dx) v |= row[dx];
    s_t[i] = v;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < dh * dw; i += K1_THREADS) {
    int ly = i / dw, lx = i - ly * dw;
    int gx = bx - elo + lx, gy = by - elo + ly;
    uint8_t v = 1;
    if (gx >= 0 && gx < W && gy >= 0 && gy < H) {
      v = 0;
      for (int dy = 0; dy < k; ++dy) v |= s_t[(ly + dy) * dw + lx];
    }
    s_d[i] = v;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < dh * K1_TW; i += K1_THREADS) {
    int ly = i / K1_TW, lx = i - ly * K1_TW;
    const uint8_t* row = s_d + ly * dw + lx;
    uint8_t v = 1;
    for (int dx = 0; dx < k; ++dx) v &= row[dx];
    s_e[i] = v;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < K1_TW * K1_TH; i += K1_THREADS) {
    int ly = i / K1_TW, lx = i - ly * K1_TW;
    int gx = bx + lx, gy = by + ly;
    if (gx >= W || gy >= H) continue;
    uint8_t v = 1;
    for (int dy = 0; dy < k; ++dy) v &= s_e[(ly + dy) * K1_TW + lx];
    out[(size_t)gy * W + gx] = v;
  }
}

// ----------------------------------------------------------------------------
// K1c: connected components by lock-free union-find (min-label roots).
// One labelling handles both classes at once: a pixel merges with a neighbour of
// the same value.  conn8_fg adds the diagonal neighbours for foreground pixels
// (OpenCV contour convention: 8-connected foreground, 4-connected background).
// Labels are slide-local pixel indices; the extra node n_s of each slide is the
// implicit zero frame around the image.
// ----------------------------------------------------------------------------
__device__ __forceinline__ int uf_find(const int* L, int x) {
  int p = L[x];
  while (p != x) {
    x = p;
    p = L[x];
  }
  return x;
}
__device__ __forceinline__ void uf_union(int* L, int a, int b) {
  bool done;
  do {
    a = uf_find(L, a);
    b = uf_find(L, b);
    if (a < b) {
      int old = atomicMin(&L[b], a);
      done = (old == b);
      b = old;
    } else if (b < a) {
      int old = atomicMin(&L[a], b);
      done = (old == a);
      a = old;
    } else {
      done = true;
    }
  } while (!done);
}

constexpr int CC_TW = 32, CC_TH = 32;   // 1024-pixel tiles, 1 thread per pixel, one warp per tile row

// ---- legacy labelling (one union per pixel and neighbour), kept for A/B (WSI_B200_SEG_LEGACY=1) ----
__global__ void __launch_bounds__(CC_TW* CC_TH)
    k_ccl_local_v1(const uint8_t* __restrict__ mask_all, SegBatch sb, int conn8_fg, int* __restrict__ L_all) {
  __shared__ int s_L[CC_TW * CC_TH];
  __shared__ uint8_t s_v[CC_TW * CC_TH];
  const int s = find_slide(sb.ctile_base, sb.n, blockIdx.x);
  const SegSlide sl = sb.slides[s];
  const int tile = (int)(blockIdx.x - sb.ctile_base[s]);
  const int ty = tile / sl.ctiles_x, tx = tile - ty * sl.ctiles_x;
  const uint8_t* __restrict__ mask = mask_all + sl.mask_off;
  int* __restrict__ L = L_all + sl.lab_off;
  const int H = sl.H, W = sl.W;
  const int lx = threadIdx.x % CC_TW, ly = threadIdx.x / CC_TW;
  const int gx = tx * CC_TW + lx, gy = ty * CC_TH + ly;
  const bool in = gx < W && gy < H;
  const int t = threadIdx.x;
  uint8_t v = in ? mask[(size_t)gy * W + gx] : 2;   // 2 = outside, never merges
  s_v[t] = v;
  s_L[t] = t;
  __syncthreads();
  if (in) {
    if (lx > 0 && s_v[t - 1] == v) uf_union(s_L, t, t - 1);
    if (ly > 0 && s_v[t - CC_TW] == v) uf_union(s_L, t, t - CC_TW);
    if (conn8_fg && v == 1 && ly > 0) {
      if (lx > 0 && s_v[t - CC_TW - 1] == 1) uf_union(s_L, t, t - CC_TW - 1);
      if (lx < CC_TW - 1 && s_v[t - CC_TW + 1] == 1) uf_union(s_L, t, t - CC_TW + 1);
    }
  }
  __syncthreads();
  if (in) {
    int rt = uf_find(s_L, t);
    int rx = rt % CC_TW, ry = rt / CC_TW;
    L[(size_t)gy * W + gx] = (ty * CC_TH + ry) * W + tx * CC_TW + rx;
  }
  if (blockIdx.x == sb.ctile_base[s] && t == 0) L[H * W] = H * W;   // frame node
}

__global__ void __launch_bounds__(256)
    k_ccl_border_v1(const uint8_t* __restrict__ mask_all, SegBatch sb, long long total, int conn8_fg, int frame,
                    int* __restrict__ L_all) {
  const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int s = find_slide_cta(sb.px_base, sb.n, g < total ? g : total - 1);
  if (g >= total) return;
  const SegSlide sl = sb.slides[s];
  const int idx = (int)(g - sb.px_base[s]);
  const uint8_t* __restrict__ mask = mask_all + sl.mask_off;
  int* __restrict__ L = L_all + sl.lab_off;
  const int H = sl.H, W = sl.W;
  const int gy = idx / W, gx = idx - gy * W;
  const bool left_edge = (gx % CC_TW) == 0, top_edge = (gy % CC_TH) == 0, right_edge = (gx % CC_TW) == CC_TW - 1;
  const bool img_border = gx == 0 || gy == 0 || gx == W - 1 || gy == H - 1;
  if (!(left_edge || top_edge || right_edge || img_border)) return;
  const uint8_t v = mask[idx];
  if (left_edge && gx > 0 && mask[idx - 1] == v) uf_union(L, idx, idx - 1);
  if (top_edge && gy > 0 && mask[idx - W] == v) uf_union(L, idx, idx - W);
  if (conn8_fg && v == 1 && gy > 0) {
    if ((top_edge || left_edge) && gx > 0 && mask[idx - W - 1] == 1) uf_union(L, idx, idx - W - 1);
    if ((top_edge || right_edge) && gx < W - 1 && mask[idx - W + 1] == 1) uf_union(L, idx, idx - W + 1);
  }
  if (frame && v == 0 && img_border) uf_union(L, idx, H * W);
}

// ---- run-based labelling ----
// Phase 1: a warp owns one 32-pixel tile row.  Ballots give the row as bit words; a pixel's first label is the
// start of its horizontal run (clz on the run-start bits), so the row needs no unions at all.  Two vertically
// adjacent runs of equal value overlap in columns [a, b] where a is the start of one of them: ONE union per pair
// of touching runs, at the lanes where either row starts a run.  8-connectivity of the foreground adds only the
// pure diagonal contacts (both shared 4-neighbours are background); every other diagonal pair is already
// connected through a horizontal run and a vertical union.  Blob-shaped masks go from ~2 atomics per pixel to a
// handful per tile.
__global__ void __launch_bounds__(CC_TW* CC_TH)
    k_ccl_local(const uint8_t* __restrict__ mask_all, SegBatch sb, int conn8_fg, int* __restrict__ L_all) {
  __shared__ int s_L[CC_TW * CC_TH];
  __shared__ unsigned s_fg[CC_TH], s_st[CC_TH];
  const int s = find_slide(sb.ctile_base, sb.n, blockIdx.x);
  const SegSlide sl = sb.slides[s];
  const int tile = (int)(blockIdx.x - sb.ctile_base[s]);
  const int ty = tile / sl.ctiles_x, tx = tile - ty * sl.ctiles_x;
  const uint8_t* __restrict__ mask = mask_all + sl.mask_off;
  int* __restrict__ L = L_all + sl.lab_off;
  const int H = sl.H, W = sl.W;
  const int t = threadIdx.x;
  const int lane = t & 31, ly = t >> 5;
  const int gx = tx * CC_TW + lane, gy = ty * CC_TH + ly;
  const bool in = gx < W && gy < H;
  const bool v = in && mask[(size_t)gy * W + gx] != 0;
  const unsigned va = __ballot_sync(0xffffffffu, in);
  const unsigned fg = __ballot_sync(0xffffffffu, v);
  const unsigned st = (fg ^ (fg << 1)) | (va ^ (va << 1)) | 1u;      // bit x: a run starts at column x
  const int rs = 31 - __clz(st & (0xffffffffu >> (31 - lane)));      // start of this lane's run
  s_L[t] = (ly << 5) + rs;
  if (lane == 0) {
    s_fg[ly] = fg;
    s_st[ly] = st;
  }
  __syncthreads();
  if (in && ly > 0) {
    // the pixel above is inside the image whenever this one is
    const unsigned fgu = s_fg[ly - 1], stu = s_st[ly - 1];
    if (((((fg ^ fgu) >> lane) & 1u) == 0u) && (((st | stu) >> lane) & 1u)) uf_union(s_L, t, t - CC_TW);
    if (conn8_fg) {
      const unsigned dl = fg & (fgu << 1) & ~fgu & ~(fg << 1);       // up-left is fg, up and left are bg
      const unsigned dr = fg & (fgu >> 1) & ~fgu & ~(fg >> 1);       // up-right is fg, up and right are bg
      if ((dl >> lane) & 1u) uf_union(s_L, t, t - CC_TW - 1);
      if ((dr >> lane) & 1u) uf_union(s_L, t, t - CC_TW + 1);
    }
  }
  __syncthreads();
  if (in) {
    const int rt = uf_find(s_L, t);
    L[(size_t)gy * W + gx] = (ty * CC_TH + (rt >> 5)) * W + tx * CC_TW + (rt & 31);
  }
  if (blockIdx.x == sb.ctile_base[s] && t == 0) L[H * W] = H * W;   // frame node
}

// Phase 2: merge across tile borders, one 128-thread CTA per tile: threads 0..31 walk the top row (unions with the
// row above, and the diagonal contacts that cross a tile border upwards), 32..63 the left column (unions with the
// column to the left, diagonal contacts that cross only the vertical border), 64..95 / 96..127 the image's right
// column / bottom row where the tile contains them (frame ties).  As in phase 1 a union is issued only where a
// run starts on either side of the border, so a large region costs one global atomic per tile edge.
__global__ void __launch_bounds__(128)
    k_ccl_border(const uint8_t* __restrict__ mask_all, SegBatch sb, int conn8_fg, int frame, int* __restrict__ L_all) {
  const int s = find_slide(sb.ctile_base, sb.n, blockIdx.x);
  const SegSlide sl = sb.slides[s];
  const int tile = (int)(blockIdx.x - sb.ctile_base[s]);
  const int ty = tile / sl.ctiles_x, tx = tile - ty * sl.ctiles_x;
  const uint8_t* __restrict__ mask = mask_all + sl.mask_off;
  int* __restrict__ L = L_all + sl.lab_off;
  const int H = sl.H, W = sl.W;
  const int bx = tx * CC_TW, by = ty * CC_TH;
  const int j = threadIdx.x & 31, role = threadIdx.x >> 5;
  if (role == 0) {                       // top row of the tile
    const int gx = bx + j, gy = by;
    if (gx >= W || gy >= H) return;
    const int idx = gy * W + gx;
    const uint8_t v = mask[idx];
    if (gy > 0) {
      const uint8_t up = mask[idx - W];
      const uint8_t lf = gx > 0 ? mask[idx - 1] : 2, ul = gx > 0 ? mask[idx - W - 1] : 2;
      if (up == v && (j == 0 || lf != v || ul != v)) uf_union(L, idx, idx - W);
      if (conn8_fg && v == 1 && up == 0) {
        if (ul == 1 && lf == 0) uf_union(L, idx, idx - W - 1);
        if (gx < W - 1 && mask[idx - W + 1] == 1 && mask[idx + 1] == 0) uf_union(L, idx, idx - W + 1);
      }
    } else if (frame && v == 0 && (j == 0 || mask[idx - 1] != 0)) {
      uf_union(L, idx, H * W);           // image top border
    }
  } else if (role == 1) {                // left column of the tile
    const int gx = bx, gy = by + j;
    if (gx >= W || gy >= H) return;
    const int idx = gy * W + gx;
    const uint8_t v = mask[idx];
    if (gx > 0) {
      const uint8_t lf = mask[idx - 1];
      const uint8_t up = gy > 0 ? mask[idx - W] : 2, ul = gy > 0 ? mask[idx - W - 1] : 2;
      if (lf == v && (j == 0 || up != v || ul != v)) uf_union(L, idx, idx - 1);
      if (conn8_fg && v == 1 && lf == 0) {
        if (j > 0 && ul == 1 && up == 0) uf_union(L, idx, idx - W - 1);
        if (j < CC_TH - 1 && gy < H - 1 && mask[idx + W - 1] == 1 && mask[idx + W] == 0) uf_union(L, idx, idx + W - 1);
      }
    } else if (frame && v == 0 && (j == 0 || mask[idx - W] != 0)) {
      uf_union(L, idx, H * W);           // image left border
    }
  } else if (role == 2) {                // image right border, where this tile holds column W-1
    if (!frame || bx + CC_TW < W) return;
    const int gx = W - 1, gy = by + j;
    if (gy >= H) return;
    const int idx = gy * W + gx;
    if (mask[idx] == 0 && (j == 0 || mask[idx - W] != 0)) uf_union(L, idx, H * W);
  } else {                               // image bottom border, where this tile holds row H-1
    if (!frame || by + CC_TH < H) return;
    const int gx = bx + j, gy = H - 1;
    if (gx >= W) return;
    const int idx = gy * W + gx;
    if (mask[idx] == 0 && (j == 0 || mask[idx - 1] != 0)) uf_union(L, idx, H * W);
  }
}

// ---- per-pixel passes: PXT pixels per thread (iteration it of a CTA handles pixel first + it * 256 + tid, coalesced) ----
constexpr int PXT = 4;
constexpr int PX_CTA = 256 * PXT;

// slide of the CTA's first pixel: one binary search per CTA (thread 0) + broadcast.  Called by ALL threads, once.
__device__ __forceinline__ int cta_slide(const long long* __restrict__ base, int n, long long first) {
  __shared__ int s_first;
  if (threadIdx.x == 0) s_first = find_slide(base, n, first);
  __syncthreads();
  return s_first;
}
__device__ __forceinline__ int advance_slide(const long long* __restrict__ base, int n, int s, long long g) {
  while (s + 1 < n && base[s + 1] <= g) ++s;
  return s;
}

// Phase 3: flatten + per-root area of class `cls` pixels (one atomic per CTA pass when all its pixels share a root,
// warp-aggregated atomics otherwise).  Transient marks 2 (fg -> bg) / 3 (bg -> fg) are folded to 0 / 1 here.
__global__ void __launch_bounds__(256)
    k_ccl_flatten_area(const uint8_t* mask_all_c, SegBatch sb, long long total, int cls,
                       int* __restrict__ L_all, int* __restrict__ area_all) {
  uint8_t* mask_all = const_cast<uint8_t*>(mask_all_c);
  const long long first = (long long)blockIdx.x * PX_CTA;
  const int s0 = cta_slide(sb.px_base, sb.n, first);
  __shared__ int s_key;
#pragma unroll 1
  for (int it = 0; it < PXT; ++it) {
    const long long g = first + it * 256 + threadIdx.x;
    const bool live = g < total;
    int root = -1 - (int)(threadIdx.x & 31);
    bool act = false;
    if (live) {
      const int s = advance_slide(sb.px_base, sb.n, s0, g);
      const SegSlide sl = sb.slides[s];
      const int idx = (int)(g - sb.px_base[s]);
      int* L = L_all + sl.lab_off;
      const int r = uf_find(L, idx);
      L[idx] = r;
      uint8_t m = (mask_all + sl.mask_off)[idx];
      if (m > 1) {
        m &= 1;
        (mask_all + sl.mask_off)[idx] = m;
      }
      act = m == cls;
      if (act) root = (int)(sl.lab_off + r);   // unique across slides inside one warp: the global label slot
    }
    if (cls < 0) continue;   // flatten only (kernel-uniform)
    __syncthreads();
    if (threadIdx.x == 0) s_key = root;
    __syncthreads();
    const bool same = act && root == s_key;
    const int c = __syncthreads_count(same);
    if (threadIdx.x == 0 && c) atomicAdd(&area_all[root], c);
    if (same) {
      act = false;
      root = -1 - (int)(threadIdx.x & 31);
    }
    const unsigned peers = __match_any_sync(0xffffffffu, root);
    if (act && (__ffs(peers) - 1) == (int)(threadIdx.x & 31)) atomicAdd(&area_all[root], __popc(peers));
  }
}

// cls==1: remove foreground components with area < thr; cls==0: fill background components with area < thr.
// mark == 0: write the new class; else write the transient mark (2: fg -> bg, 3: bg -> fg; value & 1 is the class)
// that the incremental merge kernels look for.
__global__ void __launch_bounds__(256)
    k_ccl_apply(uint8_t* __restrict__ mask_all, SegBatch sb, long long total, const int* __restrict__ L_all,
                const int* __restrict__ area_all, int cls, int mark) {
  const long long first = (long long)blockIdx.x * PX_CTA;
  const int s0 = cta_slide(sb.px_base, sb.n, first);
#pragma unroll
  for (int it = 0; it < PXT; ++it) {
    const long long g = first + it * 256 + threadIdx.x;
    if (g >= total) continue;
    const int s = advance_slide(sb.px_base, sb.n, s0, g);
    const SegSlide sl = sb.slides[s];
    const int idx = (int)(g - sb.px_base[s]);
    uint8_t* mask = mask_all + sl.mask_off;
    const int thr = cls ? sl.min_obj : sl.min_hole;
    if (thr <= 0) continue;
    if (mask[idx] == cls && area_all[sl.lab_off + L_all[sl.lab_off + idx]] < thr) mask[idx] = mark ? mark : 1 - cls;
  }
}

// Incremental relabelling: pixels that just changed class (value == mark) join the neighbouring components of their new
// class -- a handful of unions on the existing forest instead of a second labelling of the whole cohort.  Pixels
// carrying the same mark belonged to one old component and already share a root.
__global__ void __launch_bounds__(256)
    k_ccl_merge_flipped(const uint8_t* __restrict__ mask_all, SegBatch sb, long long total, int mark,
                        int* __restrict__ L_all) {
  const long long first = (long long)blockIdx.x * PX_CTA;
  const int s0 = cta_slide(sb.px_base, sb.n, first);
#pragma unroll
  for (int it = 0; it < PXT; ++it) {
    const long long g = first + it * 256 + threadIdx.x;
    if (g >= total) continue;
    const int s = advance_slide(sb.px_base, sb.n, s0, g);
    const SegSlide sl = sb.slides[s];
    const int idx = (int)(g - sb.px_base[s]);
    const uint8_t* __restrict__ mask = mask_all + sl.mask_off;
    if (mask[idx] != mark) continue;
    int* __restrict__ L = L_all + sl.lab_off;
    const int H = sl.H, W = sl.W, cls = mark & 1;
    const int gy = idx / W, gx = idx - gy * W;
    auto link = [&](int q) {
      const uint8_t m = mask[q];
      if (m != mark && (m & 1) == cls) uf_union(L, idx, q);
    };
    if (gx > 0) link(idx - 1);
    if (gx < W - 1) link(idx + 1);
    if (gy > 0) link(idx - W);
    if (gy < H - 1) link(idx + W);
  }
}

// From the 4-connected labelling of both classes to the one the outermost test needs: filled holes (mark 3) join the
// foreground around them, foreground pixels gain their pure diagonal contacts (8-connectivity), border background is
// tied to the frame node (one union per run of border background).
__global__ void __launch_bounds__(256)
    k_ccl_outer_links(const uint8_t* __restrict__ mask_all, SegBatch sb, long long total, int* __restrict__ L_all) {
  const long long first = (long long)blockIdx.x * PX_CTA;
  const int s0 = cta_slide(sb.px_base, sb.n, first);
#pragma unroll
  for (int it = 0; it < PXT; ++it) {
    const long long g = first + it * 256 + threadIdx.x;
    if (g >= total) continue;
    const int s = advance_slide(sb.px_base, sb.n, s0, g);
    const SegSlide sl = sb.slides[s];
    const int idx = (int)(g - sb.px_base[s]);
    const uint8_t* __restrict__ mask = mask_all + sl.mask_off;
    int* __restrict__ L = L_all + sl.lab_off;
    const int H = sl.H, W = sl.W;
    const int gy = idx / W, gx = idx - gy * W;
    const uint8_t m = mask[idx];
    auto fg = [&](int q) { return (mask[q] & 1) != 0; };
    if (m & 1) {
      if (m == 3) {
        if (gx > 0 && mask[idx - 1] == 1) uf_union(L, idx, idx - 1);
        if (gx < W - 1 && mask[idx + 1] == 1) uf_union(L, idx, idx + 1);
        if (gy > 0 && mask[idx - W] == 1) uf_union(L, idx, idx - W);
        if (gy < H - 1 && mask[idx + W] == 1) uf_union(L, idx, idx + W);
      }
      if (gy > 0 && !fg(idx - W)) {
        if (gx > 0 && fg(idx - W - 1) && !fg(idx - 1)) uf_union(L, idx, idx - W - 1);
        if (gx < W - 1 && fg(idx - W + 1) && !fg(idx + 1)) uf_union(L, idx, idx - W + 1);
      }
    } else {
      bool tie = false;
      if (gy == 0 || gy == H - 1) tie = gx == 0 || fg(idx - 1);
      if (!tie && (gx == 0 || gx == W - 1)) tie = gy == 0 || fg(idx - W);
      if (tie) uf_union(L, idx, H * W);
    }
  }
}

// fold the transient marks when no later kernel does it
__global__ void __launch_bounds__(256) k_mask_norm(uint8_t* __restrict__ mask, long long nbytes) {
  const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (g < nbytes && mask[g] > 1) mask[g] &= 1;
}

// outermost: a foreground component is outermost iff one of its pixels is on the image border or
// 4-adjacent to the background component tied to the frame node.  (L flattened, incl. frame node.)
__global__ void __launch_bounds__(256)
    k_outer_flag(const uint8_t* __restrict__ mask_all, SegBatch sb, long long total, const int* __restrict__ L_all,
                 int* __restrict__ flag_all) {
  const long long first = (long long)blockIdx.x * PX_CTA;
  const int s0 = cta_slide(sb.px_base, sb.n, first);
#pragma unroll
  for (int it = 0; it < PXT; ++it) {
    const long long g = first + it * 256 + threadIdx.x;
    if (g >= total) continue;
    const int s = advance_slide(sb.px_base, sb.n, s0, g);
    const SegSlide sl = sb.slides[s];
    const int idx = (int)(g - sb.px_base[s]);
    const uint8_t* __restrict__ mask = mask_all + sl.mask_off;
    if (mask[idx] != 1) continue;
    const int* __restrict__ L = L_all + sl.lab_off;
    const int H = sl.H, W = sl.W;
    const int gy = idx / W, gx = idx - gy * W;
    bool o = gx == 0 || gy == 0 || gx == W - 1 || gy == H - 1;
    if (!o) {
      // only a pixel with a background 4-neighbour can touch the frame-connected background
      const bool b0 = mask[idx - 1] == 0, b1 = mask[idx + 1] == 0, b2 = mask[idx - W] == 0, b3 = mask[idx + W] == 0;
      if (b0 || b1 || b2 || b3) {
        const int fr = uf_find(L, H * W);
        o = (b0 && L[idx - 1] == fr) || (b1 && L[idx + 1] == fr) || (b2 && L[idx - W] == fr) || (b3 && L[idx + W] == fr);
      }
    }
    if (o) flag_all[sl.lab_off + L[idx]] = 1;
  }
}

__global__ void __launch_bounds__(256)
    k_outer_apply(const uint8_t* __restrict__ mask_all, SegBatch sb, long long total, const int* __restrict__ L_all,
                  const int* __restrict__ flag_all, uint8_t* __restrict__ outer_all, int* __restrict__ info) {
  const long long first = (long long)blockIdx.x * PX_CTA;
  const int s0 = cta_slide(sb.px_base, sb.n, first);
#pragma unroll
  for (int it = 0; it < PXT; ++it) {
    const long long g = first + it * 256 + threadIdx.x;
    if (g >= total) continue;
    const int s = advance_slide(sb.px_base, sb.n, s0, g);
    const SegSlide sl = sb.slides[s];
    const int idx = (int)(g - sb.px_base[s]);
    const uint8_t m = (mask_all + sl.mask_off)[idx];
    (outer_all + sl.mask_off)[idx] = (m == 1 && flag_all[sl.lab_off + L_all[sl.lab_off + idx]]) ? 1 : 0;
    // per-slide "the mask has tissue" flag (replaces a host-side mask.any()): the store of a slide's FIRST tissue
    // pixel in this thread's stripe is enough -- benign same-value stores, no warp collectives
    if (info && m == 1 && (idx == 0 || (mask_all + sl.mask_off)[idx - 1] != 1)) info[4 * s + 1] = 1;
  }
}

// ----------------------------------------------------------------------------
// K2: patch-grid filter + order-preserving compaction
// ----------------------------------------------------------------------------
struct GridArgs {
  const uint8_t* outer;
  int Ht, Wt;
  long long W, H;
  int half, step, ni, nj;
};

__device__ __forceinline__ int outer_at(const GridArgs& g, long long i, long long j) {
  return (i >= 0 && i < g.Ht && j >= 0 && j < g.Wt) ? g.outer[i * g.Wt + j] : 0;
}

// exact integer point-in-region test (oracle/patchgrid.py::raster_rule)
__device__ __forceinline__ bool grid_inside(const GridArgs& g, int cand) {
  const int ci = cand / g.nj, cj = cand - ci * g.nj;
  const long long cx = g.half + (long long)ci * g.step;
  const long long cy = g.half + (long long)cj * g.step;
  const long long nu = cx * g.Wt, nv = cy * g.Ht;
  const long long j = nu / g.W, i = nv / g.H;
  const long long ru = nu - j * g.W, rv = nv - i * g.H;
  const int a = outer_at(g, i, j), b = outer_at(g, i, j + 1), c = outer_at(g, i + 1, j),
            d = outer_at(g, i + 1, j + 1);
  if (ru > 0 && rv > 0) {
    const int cnt = a + b + c + d;
    if (cnt == 4) return true;
    if (cnt != 3) return false;
    const long long lu = ru * g.H, lv = rv * g.W, wh = g.W * g.H;
    if (!a) return lu + lv > wh;
    if (!b) return lv > lu;
    if (!c) return lu > lv;
    return lu + lv < wh;
  }
  if (ru == 0 && rv > 0)
    return a && c && (b || d) && (outer_at(g, i, j - 1) || outer_at(g, i + 1, j - 1));
  if (ru > 0 && rv == 0)
    return a && b && (c || d) && (outer_at(g, i - 1, j) || outer_at(g, i - 1, j + 1));
  return a && b && c && outer_at(g, i - 1, j) && outer_at(g, i, j - 1);
}

constexpr int GRID_THREADS = 256;

__global__ void __launch_bounds__(GRID_THREADS) k_grid_count(GridArgs g, int ncand, int* __restrict__ blk) {
  __shared__ int s_w[GRID_THREADS / 32];
  int cand = blockIdx.x * GRID_THREADS + threadIdx.x;
  bool in = cand < ncand && grid_inside(g, cand);
  unsigned bal = __ballot_sync(0xffffffffu, in);
  if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = __popc(bal);
  __syncthreads();
  if (threadIdx.x == 0) {
    int s = 0;
    for (int w = 0; w < GRID_THREADS / 32; ++w) s += s_w[w];
    blk[blockIdx.x] = s;
  }
}

// single-block exclusive scan over the per-block counts; total -> *count
__global__ void __launch_bounds__(1024) k_grid_scan(int* __restrict__ blk, int nblk, int* __restrict__ count) {
  __shared__ int s_part[1024];
  const int per = (nblk + 1023) / 1024;
  const int b0 = min(threadIdx.x * per, nblk), b1 = min(b0 + per, nblk);
  int s = 0;
  for (int i = b0; i < b1; ++i) s += blk[i];
  s_part[threadIdx.x] = s;
  __syncthreads();
  for (int off = 1; off < 1024; off <<= 1) {   // Hillis-Steele inclusive scan
    int v = threadIdx.x >= off ? s_part[threadIdx.x - off] : 0;
    __syncthreads();
    s_part[threadIdx.x] += v;
    __syncthreads();
  }
  int run = s_part[threadIdx.x] - s;
  for (int i = b0; i < b1; ++i) {
    int c = blk[i];
    blk[i] = run;
    run += c;
  }
  if (threadIdx.x == 1023) *count = s_part[1023];
}

__global__ void __launch_bounds__(GRID_THREADS)
    k_grid_write(GridArgs g, int ncand, const int* __restrict__ blk, int capacity, int32_t* __restrict__ coords) {
  __shared__ int s_w[GRID_THREADS / 32];
  int cand = blockIdx.x * GRID_THREADS + threadIdx.x;
  bool in = cand < ncand && grid_inside(g, cand);
  unsigned bal = __ballot_sync(0xffffffffu, in);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) s_w[warp] = __popc(bal);
  __syncthreads();
  int base = blk[blockIdx.x];
  for (int w = 0; w < warp; ++w) base += s_w[w];
  if (in) {
    int pos = base + __popc(bal & ((1u << lane) - 1u));
    if (pos < capacity) {
      const int ci = cand / g.nj, cj = cand - ci * g.nj;
      int2 xy;
      xy.x = ci * g.step;   // (half + ci*step) - half
      xy.y = cj * g.step;
      reinterpret_cast<int2*>(coords)[pos] = xy;
    }
  }
}

bool g_sdiv_ready = false;

int upload_sdiv() {
  if (g_sdiv_ready) return WSI_OK;
  int h[256];
  h[0] = 0;
  for (int i = 1; i < 256; ++i) {
    double v = (double)(255 << 12) / (double)i;
    h[i] = (int)__builtin_rint(v);
  }
  WSI_CUDA_CHECK(cudaMemcpyToSymbol(c_sdiv, h, sizeof(h)));
  g_sdiv_ready = true;
  return WSI_OK;
}

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

struct BatchLayout {
  size_t off_desc, off_tile, off_ctile, off_px, off_hist, off_thr, off_info, off_tmp, off_L, off_aux, total;
  size_t head_bytes;      // descriptor tables: one host->device copy
  long long npx, ntiles, nctiles;
};

BatchLayout batch_layout(int n, const int* H, const int* W) {
  BatchLayout b;
  b.npx = 0;
  b.ntiles = 0;
  b.nctiles = 0;
  for (int s = 0; s < n; ++s) {
    b.npx += (long long)H[s] * W[s];
    b.ntiles += (long long)wsi_cdiv(W[s], K1_TW) * wsi_cdiv(H[s], K1_TH);
    b.nctiles += (long long)wsi_cdiv(W[s], CC_TW) * wsi_cdiv(H[s], CC_TH);
  }
  size_t o = 0;
  b.off_desc = o;  o += align_up(sizeof(SegSlide) * n, 256);
  b.off_tile = o;  o += align_up(sizeof(long long) * (n + 1), 256);
  b.off_ctile = o; o += align_up(sizeof(long long) * (n + 1), 256);
  b.off_px = o;    o += align_up(sizeof(long long) * (n + 1), 256);
  b.head_bytes = o;
  b.off_hist = o;  o += align_up(sizeof(int) * 256 * (size_t)n, 256);     // Otsu mode: per-slide histogram of S_med
  b.off_thr = o;   o += align_up(sizeof(int) * (size_t)n, 256);           //            per-slide threshold
  b.off_info = o;  o += align_up(sizeof(int) * 4 * (size_t)n, 256);       // per-slide {threshold, has_tissue, 0, 0}
  b.off_tmp = o;   o += align_up((size_t)b.npx, 256);
  b.off_L = o;     o += align_up(((size_t)b.npx + n) * sizeof(int), 256);
  b.off_aux = o;   o += align_up(((size_t)b.npx + n) * sizeof(int), 256);
  b.total = o;
  return b;
}

bool seg_legacy() {
  const char* e = getenv("WSI_B200_SEG_LEGACY");
  return e && e[0] == '1';
}

__global__ void k_info_thr(const int* __restrict__ thr, int fixed_thr, int n, int* __restrict__ info) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s < n) info[4 * s] = thr ? thr[s] : fixed_thr;
}

}  // namespace

extern "C" {

size_t wsi_seg_batch_workspace_bytes(int n, const int* h_H, const int* h_W) {
  if (n <= 0 || !h_H || !h_W) return 0;
  return batch_layout(n, h_H, h_W).total;
}

size_t wsi_seg_workspace_bytes(int H, int W) { return wsi_seg_batch_workspace_bytes(1, &H, &W); }

static int seg_prepare(int n, const int* h_H, const int* h_W, const int* h_pitch, const long long* h_rgb_off,
                       const long long* h_mask_off, const int* h_min_object_px, const int* h_min_hole_px, void* d_ws,
                       size_t ws_bytes, cudaStream_t stream, SegBatch* sb_out, BatchLayout* bl_out) {
  WSI_ARG_CHECK(n > 0, "wsi_seg_tissue: empty batch");
  for (int s = 0; s < n; ++s) {
    WSI_ARG_CHECK(h_H[s] > 0 && h_W[s] > 0, "wsi_seg_tissue: bad geometry %dx%d", h_H[s], h_W[s]);
    WSI_ARG_CHECK((long long)h_H[s] * h_W[s] < (1ll << 30), "wsi_seg_tissue: thumbnail too large");
  }
  const BatchLayout bl = batch_layout(n, h_H, h_W);
  WSI_ARG_CHECK(ws_bytes >= bl.total, "wsi_seg_tissue: workspace too small (%zu < %zu)", ws_bytes, bl.total);
  WSI_ARG_CHECK(bl.npx + n < (1ll << 31), "wsi_seg_tissue: cohort too large for one call");
  // descriptor tables, staged contiguously on the host in the layout of the workspace head: ONE copy
  std::vector<uint8_t> head(bl.head_bytes, 0);
  SegSlide* hs = (SegSlide*)(head.data() + bl.off_desc);
  long long* tb = (long long*)(head.data() + bl.off_tile);
  long long* cb = (long long*)(head.data() + bl.off_ctile);
  long long* pb = (long long*)(head.data() + bl.off_px);
  tb[0] = cb[0] = pb[0] = 0;
  long long max_mask_end = 0;
  for (int s = 0; s < n; ++s) {
    SegSlide& d = hs[s];
    d.H = h_H[s]; d.W = h_W[s]; d.pitch = h_pitch ? h_pitch[s] : 0;
    d.min_obj = h_min_object_px ? h_min_object_px[s] : 0;
    d.min_hole = h_min_hole_px ? h_min_hole_px[s] : 0;
    d.tiles_x = wsi_cdiv(d.W, K1_TW);
    d.ctiles_x = wsi_cdiv(d.W, CC_TW);
    d.pad_ = 0;
    d.rgb_off = h_rgb_off ? h_rgb_off[s] : 0;
    d.mask_off = h_mask_off[s];
    d.lab_off = pb[s] + s;
    tb[s + 1] = tb[s] + (long long)d.tiles_x * wsi_cdiv(d.H, K1_TH);
    cb[s + 1] = cb[s] + (long long)d.ctiles_x * wsi_cdiv(d.H, CC_TH);
    pb[s + 1] = pb[s] + (long long)d.H * d.W;
    WSI_ARG_CHECK(h_mask_off[s] >= 0, "wsi_seg_tissue: negative mask offset");
    const long long e = h_mask_off[s] + (long long)d.H * d.W;
    if (e > max_mask_end) max_mask_end = e;
  }
  // the median scratch is addressed with the mask offsets, so masks must pack the cohort densely
  WSI_ARG_CHECK(max_mask_end <= bl.npx, "wsi_seg_tissue: mask offsets must pack the cohort densely");
  uint8_t* ws = (uint8_t*)d_ws;
  // pageable source: the runtime stages the bytes before returning, so `head` may go out of scope
  WSI_CUDA_CHECK(cudaMemcpyAsync(ws, head.data(), bl.head_bytes, cudaMemcpyHostToDevice, stream));
  sb_out->slides = (const SegSlide*)(ws + bl.off_desc);
  sb_out->tile_base = (const long long*)(ws + bl.off_tile);
  sb_out->ctile_base = (const long long*)(ws + bl.off_ctile);
  sb_out->px_base = (const long long*)(ws + bl.off_px);
  sb_out->n = n;
  *bl_out = bl;
  return WSI_OK;
}

// one labelling of the whole cohort: tile-local pass + border merge (+ frame ties)
static void seg_label(const SegBatch& sb, const BatchLayout& bl, const uint8_t* d_mask, int conn8_fg, int frame, int* L,
                      cudaStream_t stream) {
  if (seg_legacy()) {
    k_ccl_local_v1<<<(unsigned)bl.nctiles, CC_TW * CC_TH, 0, stream>>>(d_mask, sb, conn8_fg, L);
    k_ccl_border_v1<<<wsi_cdiv(bl.npx, 256), 256, 0, stream>>>(d_mask, sb, bl.npx, conn8_fg, frame, L);
  } else {
    k_ccl_local<<<(unsigned)bl.nctiles, CC_TW * CC_TH, 0, stream>>>(d_mask, sb, conn8_fg, L);
    k_ccl_border<<<(unsigned)bl.nctiles, 128, 0, stream>>>(d_mask, sb, conn8_fg, frame, L);
  }
  wsi_count_launch(2);
}

// 8-connected foreground labelling + frame adjacency -> outermost-component mask.
// have_base: L already holds the 4-connected labelling of both classes of the mask as it was before the pixels now
// carrying the marks 2 / 3 changed class (those with mark 2 are already merged): add the missing links instead of
// labelling again.
static int seg_outermost(const SegBatch& sb, const BatchLayout& bl, uint8_t* d_mask, uint8_t* d_outer, void* d_ws,
                         int* d_info, bool have_base, cudaStream_t stream) {
  uint8_t* ws = (uint8_t*)d_ws;
  int* L = (int*)(ws + bl.off_L);
  int* aux = (int*)(ws + bl.off_aux);
  const long long total = bl.npx;
  const size_t lab_bytes = ((size_t)bl.npx + sb.n) * sizeof(int);
  const int nb = wsi_cdiv(total, 256), nbp = wsi_cdiv(total, PX_CTA);
  (void)nb;
  if (have_base) {
    k_ccl_outer_links<<<nbp, 256, 0, stream>>>(d_mask, sb, total, L);
    wsi_count_launch(1);
  } else {
    seg_label(sb, bl, d_mask, 1, 1, L, stream);
  }
  k_ccl_flatten_area<<<nbp, 256, 0, stream>>>(d_mask, sb, total, -1, L, aux);
  WSI_CUDA_CHECK(cudaMemsetAsync(aux, 0, lab_bytes, stream));
  k_outer_flag<<<nbp, 256, 0, stream>>>(d_mask, sb, total, L, aux);
  k_outer_apply<<<nbp, 256, 0, stream>>>(d_mask, sb, total, L, aux, d_outer, d_info);
  wsi_count_launch(3);
  return WSI_OK;
}

int wsi_seg_tissue_batch_ex(const uint8_t* d_rgb, int n, const int* h_H, const int* h_W, const int* h_pitch,
                            const long long* h_rgb_off, const long long* h_mask_off, int median_filter_size,
                            int binary_threshold, int closing_kernel_size, int closing_dilation_origin,
                            int closing_erosion_origin, const int* h_min_object_px, const int* h_min_hole_px,
                            uint8_t* d_mask, uint8_t* d_outer, int* d_info, void* d_ws, size_t ws_bytes,
                            void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  WSI_ARG_CHECK(d_rgb && d_mask && d_ws && h_H && h_W && h_pitch && h_rgb_off && h_mask_off && h_min_object_px &&
                    h_min_hole_px,
                "wsi_seg_tissue: null pointer");
  // same message as the reference's ValueError (segment.py:68-72)
  WSI_ARG_CHECK(median_filter_size > 1 && (median_filter_size % 2) == 1,
                "median_filter_size must be greater than 1 and odd, but got %d", median_filter_size);
  WSI_ARG_CHECK(median_filter_size <= 31, "median_filter_size > 31 unsupported");
  WSI_ARG_CHECK(closing_kernel_size >= 1 && closing_kernel_size <= 32, "closing_kernel_size out of range");
  WSI_ARG_CHECK(binary_threshold >= WSI_SEG_THRESHOLD_OTSU, "binary_threshold must be >= 0 or WSI_SEG_THRESHOLD_OTSU");
  const int k = closing_kernel_size;
  const int dlo = closing_dilation_origin < 0 ? (k - 1) / 2 : closing_dilation_origin;
  const int elo = closing_erosion_origin < 0 ? k / 2 : closing_erosion_origin;
  WSI_ARG_CHECK(dlo < k && elo < k, "closing origins (%d, %d) must lie inside the %dx%d footprint", dlo, elo, k, k);
  for (int s = 0; s < n; ++s)
    WSI_ARG_CHECK(h_pitch[s] >= 3 * h_W[s], "wsi_seg_tissue: pitch %d < 3*W (%d)", h_pitch[s], h_W[s]);
  int rc = upload_sdiv();
  if (rc) return rc;
  SegBatch sb;
  BatchLayout bl;
  rc = seg_prepare(n, h_H, h_W, h_pitch, h_rgb_off, h_mask_off, h_min_object_px, h_min_hole_px, d_ws, ws_bytes, stream,
                   &sb, &bl);
  if (rc) return rc;
  uint8_t* ws = (uint8_t*)d_ws;
  uint8_t* tmp = ws + bl.off_tmp;
  int* L = (int*)(ws + bl.off_L);
  int* aux = (int*)(ws + bl.off_aux);
  int* hist = (int*)(ws + bl.off_hist);
  int* thr = (int*)(ws + bl.off_thr);
  int* info = d_info ? d_info : (int*)(ws + bl.off_info);
  const long long total = bl.npx;
  const size_t lab_bytes = ((size_t)bl.npx + n) * sizeof(int);
  const int nb = wsi_cdiv(total, 256), nbp = wsi_cdiv(total, PX_CTA);
  (void)nb;
  const bool legacy = seg_legacy();
  int launches = 0;
  WSI_CUDA_CHECK(cudaMemsetAsync(info, 0, sizeof(int) * 4 * (size_t)n, stream));
  {
    const int r = median_filter_size / 2;
    const size_t sw = K1_TW + 2 * r, sh = K1_TH + 2 * r;
    if (binary_threshold == WSI_SEG_THRESHOLD_OTSU) {
      // S_med values -> d_mask (scratch), histogram -> Otsu threshold per slide -> tmp = S_med > t
      WSI_CUDA_CHECK(cudaMemsetAsync(hist, 0, sizeof(int) * 256 * (size_t)n, stream));
      k_sat_median_value<<<(unsigned)bl.ntiles, K1_THREADS, sw * sh, stream>>>(d_rgb, sb, median_filter_size, d_mask,
                                                                               hist);
      k_otsu<<<wsi_cdiv(n, 32), 32, 0, stream>>>(hist, sb, thr);
      k_thresh_slide<<<nb, 256, 0, stream>>>(d_mask, sb, total, thr, tmp);
      k_info_thr<<<wsi_cdiv(n, 64), 64, 0, stream>>>(thr, 0, n, info);
      launches += 4;
    } else {
      if (legacy)
        k_sat_median_thresh_v1<<<(unsigned)bl.ntiles, K1_THREADS, sw * sh, stream>>>(d_rgb, sb, median_filter_size,
                                                                                     binary_threshold, tmp);
      else
        k_sat_median_thresh<<<(unsigned)bl.ntiles, K1_THREADS, sw * sh + sh * K1_TW, stream>>>(
            d_rgb, sb, median_filter_size, binary_threshold, tmp);
      k_info_thr<<<wsi_cdiv(n, 64), 64, 0, stream>>>(nullptr, binary_threshold, n, info);
      launches += 2;
    }
  }
  {
    const int halo = k - 1;
    const size_t iw = K1_TW + 2 * halo, ih = K1_TH + 2 * halo, dw = K1_TW + halo, dh = K1_TH + halo;
    if (legacy)
      k_closing_v1<<<(unsigned)bl.ntiles, K1_THREADS, iw * ih + dw * dh, stream>>>(tmp, sb, k, dlo, elo, d_mask);
    else
      k_closing<<<(unsigned)bl.ntiles, K1_THREADS, iw * ih + ih * dw + dh * dw + dh * K1_TW, stream>>>(tmp, sb, k, dlo,
                                                                                                    elo, d_mask);
    ++launches;
  }
  // remove_small_objects (4-conn foreground, area < min) then remove_small_holes (4-conn background)
  bool any_obj = false, any_hole = false;
  for (int s = 0; s < n; ++s) {
    any_obj = any_obj || h_min_object_px[s] > 0;
    any_hole = any_hole || h_min_hole_px[s] > 0;
  }
  bool have_base = false;
  if (legacy) {
    for (int pass = 0; pass < 2; ++pass) {
      const int cls = pass == 0 ? 1 : 0;
      if (!(cls ? any_obj : any_hole)) continue;
      seg_label(sb, bl, d_mask, 0, 0, L, stream);
      WSI_CUDA_CHECK(cudaMemsetAsync(aux, 0, lab_bytes, stream));
      k_ccl_flatten_area<<<nbp, 256, 0, stream>>>(d_mask, sb, total, cls, L, aux);
      k_ccl_apply<<<nbp, 256, 0, stream>>>(d_mask, sb, total, L, aux, cls, 0);
      launches += 2;
    }
  } else if (any_obj || any_hole || d_outer) {
    // ONE labelling of both classes; every later stage adds the few unions its class changes call for
    seg_label(sb, bl, d_mask, 0, 0, L, stream);
    have_base = true;
    bool marks = false;
    if (any_obj) {
      WSI_CUDA_CHECK(cudaMemsetAsync(aux, 0, lab_bytes, stream));
      k_ccl_flatten_area<<<nbp, 256, 0, stream>>>(d_mask, sb, total, 1, L, aux);
      k_ccl_apply<<<nbp, 256, 0, stream>>>(d_mask, sb, total, L, aux, 1, 2);
      k_ccl_merge_flipped<<<nbp, 256, 0, stream>>>(d_mask, sb, total, 2, L);
      launches += 3;
      marks = true;
    }
    if (any_hole) {
      WSI_CUDA_CHECK(cudaMemsetAsync(aux, 0, lab_bytes, stream));
      k_ccl_flatten_area<<<nbp, 256, 0, stream>>>(d_mask, sb, total, 0, L, aux);     // folds the marks 2 -> 0
      k_ccl_apply<<<nbp, 256, 0, stream>>>(d_mask, sb, total, L, aux, 0, 3);
      launches += 2;
      marks = true;
    }
    if (!d_outer && marks) {
      k_mask_norm<<<wsi_cdiv(total, 256), 256, 0, stream>>>(d_mask, total);
      ++launches;
    }
  }
  wsi_count_launch(launches);
  if (d_outer) {
    rc = seg_outermost(sb, bl, d_mask, d_outer, d_ws, info, have_base, stream);
    if (rc) return rc;
  }
  WSI_CUDA_CHECK(cudaGetLastError());
  return WSI_OK;
}

int wsi_seg_tissue_batch_u8(const uint8_t* d_rgb, int n, const int* h_H, const int* h_W, const int* h_pitch,
                            const long long* h_rgb_off, const long long* h_mask_off, int median_filter_size,
                            int binary_threshold, int closing_kernel_size, const int* h_min_object_px,
                            const int* h_min_hole_px, uint8_t* d_mask, uint8_t* d_outer, void* d_ws, size_t ws_bytes,
                            void* stream_) {
  WSI_ARG_CHECK(binary_threshold >= 0, "wsi_seg_tissue: binary_threshold must be >= 0 (Otsu: wsi_seg_tissue_batch_ex)");
  return wsi_seg_tissue_batch_ex(d_rgb, n, h_H, h_W, h_pitch, h_rgb_off, h_mask_off, median_filter_size,
                                 binary_threshold, closing_kernel_size, -1, -1, h_min_object_px, h_min_hole_px, d_mask,
                                 d_outer, nullptr, d_ws, ws_bytes, stream_);
}

int wsi_seg_tissue_u8(const uint8_t* d_rgb, int H, int W, int pitch_bytes, int median_filter_size,
                      int binary_threshold, int closing_kernel_size, int min_object_size_px,
                      int min_hole_size_px, uint8_t* d_mask, uint8_t* d_outer, void* d_ws, size_t ws_bytes,
                      void* stream_) {
  const long long zero = 0;
  return wsi_seg_tissue_batch_u8(d_rgb, 1, &H, &W, &pitch_bytes, &zero, &zero, median_filter_size, binary_threshold,
                                 closing_kernel_size, &min_object_size_px, &min_hole_size_px, d_mask, d_outer, d_ws,
                                 ws_bytes, stream_);
}

int wsi_outermost_u8(const uint8_t* d_mask, int H, int W, uint8_t* d_outer, void* d_ws, size_t ws_bytes,
                     void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  WSI_ARG_CHECK(d_mask && d_outer && d_ws, "wsi_outermost_u8: null pointer");
  const long long zero = 0;
  SegBatch sb;
  BatchLayout bl;
  int rc = seg_prepare(1, &H, &W, nullptr, nullptr, &zero, nullptr, nullptr, d_ws, ws_bytes, stream, &sb, &bl);
  if (rc) return rc;
  rc = seg_outermost(sb, bl, const_cast<uint8_t*>(d_mask), d_outer, d_ws, nullptr, false, stream);
  if (rc) return rc;
  WSI_CUDA_CHECK(cudaGetLastError());
  return WSI_OK;
}

// Generic labelling of ONE 0/1 image for other modules (csrc/cellpost.cu): min-index union-find labels of both classes
// (4-connected; conn8_fg adds the diagonals for the foreground; frame ties border background to the node H*W) and,
// for area_cls in {0, 1}, the pixel count of every component of that class at its root.  *d_L_out / *d_area_out
// point INTO the workspace (H*W + 1 ints each) and stay valid until the workspace is reused.
int wsi_ccl_label_u8(const uint8_t* d_mask, int H, int W, int conn8_fg, int frame, int area_cls, int** d_L_out,
                     int** d_area_out, void* d_ws, size_t ws_bytes, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  WSI_ARG_CHECK(d_mask && d_ws && d_L_out, "wsi_ccl_label_u8: null pointer");
  WSI_ARG_CHECK(area_cls >= -1 && area_cls <= 1, "wsi_ccl_label_u8: area_cls must be -1, 0 or 1");
  const long long zero = 0;
  SegBatch sb;
  BatchLayout bl;
  int rc = seg_prepare(1, &H, &W, nullptr, nullptr, &zero, nullptr, nullptr, d_ws, ws_bytes, stream, &sb, &bl);
  if (rc) return rc;
  uint8_t* ws = (uint8_t*)d_ws;
  int* L = (int*)(ws + bl.off_L);
  int* aux = (int*)(ws + bl.off_aux);
  seg_label(sb, bl, d_mask, conn8_fg, frame, L, stream);
  if (area_cls >= 0) WSI_CUDA_CHECK(cudaMemsetAsync(aux, 0, ((size_t)bl.npx + 1) * sizeof(int), stream));
  k_ccl_flatten_area<<<wsi_cdiv(bl.npx, PX_CTA), 256, 0, stream>>>(d_mask, sb, bl.npx, area_cls, L, aux);
  wsi_count_launch(1);
  *d_L_out = L;
  if (d_area_out) *d_area_out = aux;
  WSI_CUDA_CHECK(cudaGetLastError());
  return WSI_OK;
}

// ---- allocation-free patch grid (no counterpart in the reference) -------------------------------------------------
// Scratch of wsi_patch_grid_static(): per-block counts / offsets of k_grid_count / k_grid_scan / k_grid_write for up to
// WSI_PATCH_GRID_STATIC_MAX_CANDIDATES candidate centroids.  256 KB of device memory on every GPU that loads the
// library, held until the process exits.  In HOST code `g_grid_static` is only a stand-in for the symbol: the device
// address comes from cudaGetSymbolAddress (grid_static_ptr below), once per device with that device current.
constexpr int GRID_STATIC_MAX_BLOCKS = WSI_PATCH_GRID_STATIC_MAX_CANDIDATES / GRID_THREADS;
static __device__ __align__(256) int g_grid_static[GRID_STATIC_MAX_BLOCKS + 64];

static int grid_static_ptr(int** out) {
  constexpr int MAX_DEV = 64;
  static std::mutex mu;
  static int* ptr_of[MAX_DEV] = {};
  int dev = -1;
  WSI_CUDA_CHECK(cudaGetDevice(&dev));
  WSI_ARG_CHECK(dev >= 0 && dev < MAX_DEV, "wsi_patch_grid_static: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (!ptr_of[dev]) {
    int* p = nullptr;
    WSI_CUDA_CHECK(cudaGetSymbolAddress((void**)&p, g_grid_static));
    WSI_ARG_CHECK(p != nullptr, "wsi_patch_grid_static: no device address for the scratch array");
    ptr_of[dev] = p;
  }
  *out = ptr_of[dev];
  return WSI_OK;
}

size_t wsi_patch_grid_workspace_bytes(long long slide_w, long long slide_h, int half, int step) {
  if (step <= 0) return 0;
  long long ni = slide_w > half ? (slide_w - half + step - 1) / step : 0;
  long long nj = slide_h > half ? (slide_h - half + step - 1) / step : 0;
  long long nblk = (ni * nj + GRID_THREADS - 1) / GRID_THREADS;
  return (size_t)(nblk + 1) * sizeof(int) + 256;
}

int wsi_patch_grid(const uint8_t* d_outer, int Ht, int Wt, long long slide_w, long long slide_h, int half,
                   int step, int32_t* d_coords_xy, int capacity, int* d_count, int* h_tile_dim, void* d_ws,
                   size_t ws_bytes, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  WSI_ARG_CHECK(d_outer && d_coords_xy && d_count && d_ws, "wsi_patch_grid: null pointer");
  WSI_ARG_CHECK(Ht > 0 && Wt > 0 && slide_w > 0 && slide_h > 0, "wsi_patch_grid: bad geometry");
  WSI_ARG_CHECK(step > 0, "wsi_patch_grid: step must be positive (overlap must be in (-inf, 1))");
  WSI_ARG_CHECK(half >= 0, "wsi_patch_grid: negative half patch size");
  // exactness bound for the int64 cross products: W*H, ru*H, rv*W < 2^62
  WSI_ARG_CHECK(slide_w < (1ll << 30) && slide_h < (1ll << 30), "wsi_patch_grid: slide too large");
  long long ni = slide_w > half ? (slide_w - half + step - 1) / step : 0;   // len(range(half, W, step))
  long long nj = slide_h > half ? (slide_h - half + step - 1) / step : 0;
  WSI_ARG_CHECK(ni * nj < (1ll << 31) - GRID_THREADS, "wsi_patch_grid: too many candidates");
  WSI_ARG_CHECK((ni - 1) * step < (1ll << 31) && (nj - 1) * step < (1ll << 31), "wsi_patch_grid: int32 overflow");
  if (h_tile_dim) {
    h_tile_dim[0] = (int)ni;
    h_tile_dim[1] = (int)nj;
  }
  const int ncand = (int)(ni * nj);
  if (ncand == 0) {
    WSI_CUDA_CHECK(cudaMemsetAsync(d_count, 0, sizeof(int), stream));
    return WSI_OK;
  }
  const int nblk = wsi_cdiv(ncand, GRID_THREADS);
  WSI_ARG_CHECK(ws_bytes >= (size_t)(nblk + 1) * sizeof(int), "wsi_patch_grid: workspace too small");
  GridArgs g{d_outer, Ht, Wt, slide_w, slide_h, half, step, (int)ni, (int)nj};
  int* blk = (int*)d_ws;
  k_grid_count<<<nblk, GRID_THREADS, 0, stream>>>(g, ncand, blk);
  k_grid_scan<<<1, 1024, 0, stream>>>(blk, nblk, d_count);
  k_grid_write<<<nblk, GRID_THREADS, 0, stream>>>(g, ncand, blk, capacity, d_coords_xy);
  wsi_count_launch(3);
  WSI_CUDA_CHECK(cudaGetLastError());
  return WSI_OK;
}

// wsi_patch_grid() with its scratch in the file-scope device array above: no cudaMalloc / cudaFree / allocating call on
// this path, for callers that cannot allocate (CUDA graph capture, tight serving loops).  Same kernels, same results.
// Every call on a device shares the one array: calls that overlap on the GPU (different streams) race; calls on one
// stream are ordered.  Nothing in the array has to start at zero -- k_grid_count overwrites every entry it later reads.
// The first call on a device resolves the array's address (cudaGetSymbolAddress); make it before a graph capture.
int wsi_patch_grid_static(const uint8_t* d_outer, int Ht, int Wt, long long slide_w, long long slide_h, int half,
                          int step, int32_t* d_coords_xy, int capacity, int* d_count, int* h_tile_dim, void* stream_) {
  WSI_ARG_CHECK(step > 0, "wsi_patch_grid: step must be positive (overlap must be in (-inf, 1))");
  WSI_ARG_CHECK(half >= 0 && slide_w > 0 && slide_h > 0 && slide_w < (1ll << 30) && slide_h < (1ll << 30),
                "wsi_patch_grid_static: bad geometry");
  const long long ni = slide_w > half ? (slide_w - half + step - 1) / step : 0;
  const long long nj = slide_h > half ? (slide_h - half + step - 1) / step : 0;
  WSI_ARG_CHECK(ni * nj <= (long long)WSI_PATCH_GRID_STATIC_MAX_CANDIDATES,
                "wsi_patch_grid_static: %lld candidate centroids exceed the static scratch (max %d); use wsi_patch_grid "
                "with a caller-provided workspace",
                ni * nj, WSI_PATCH_GRID_STATIC_MAX_CANDIDATES);
  int* scratch = nullptr;
  int rc = grid_static_ptr(&scratch);
  if (rc) return rc;
  return wsi_patch_grid(d_outer, Ht, Wt, slide_w, slide_h, half, step, d_coords_xy, capacity, d_count, h_tile_dim,
                        scratch, sizeof(int) * (size_t)(GRID_STATIC_MAX_BLOCKS + 64), stream_);
}

}  // extern "C"
