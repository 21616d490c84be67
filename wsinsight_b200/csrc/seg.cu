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

// ----------------------------------------------------------------------------
// K1a: RGB -> HSV saturation (OpenCV 8-bit formula) -> (S > thresh) -> k x k box
// count with replicate border >= (k*k+1)/2   [== medianBlur(S,k) > thresh]
// ----------------------------------------------------------------------------
__constant__ int c_sdiv[256];

constexpr int K1_TW = 64;
constexpr int K1_TH = 16;
constexpr int K1_THREADS = 256;

__global__ void __launch_bounds__(K1_THREADS)
    k_sat_median_thresh(const uint8_t* __restrict__ rgb_all, SegBatch sb, int k, int thresh,
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
    const uint8_t* p = rgb + (size_t)gy * sl.pitch + (size_t)gx * 3;
    int R = p[0], G = p[1], B = p[2];
    int v = max(R, max(G, B));
    int d = v - min(R, min(G, B));
    int sat = (d * c_sdiv[v] + (1 << 11)) >> 12;
    s_b[i] = (sat > thresh) ? 1 : 0;
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

// ----------------------------------------------------------------------------
// K1b: closing with ones(k,k):  dilation offsets [-(k-1)/2 .. k/2] (outside = 0)
// then erosion offsets [-(k/2) .. (k-1)/2] (outside = 1)  (scipy / skimage convention)
// ----------------------------------------------------------------------------
__global__ void __launch_bounds__(K1_THREADS)
    k_closing(const uint8_t* __restrict__ in_all, SegBatch sb, int k, uint8_t* __restrict__ out_all) {
  extern __shared__ uint8_t s_m[];
  const int s = find_slide(sb.tile_base, sb.n, blockIdx.x);
  const SegSlide sl = sb.slides[s];
  const int tile = (int)(blockIdx.x - sb.tile_base[s]);
  const int ty = tile / sl.tiles_x, tx = tile - ty * sl.tiles_x;
  const uint8_t* __restrict__ in = in_all + sl.mask_off;
  uint8_t* __restrict__ out = out_all + sl.mask_off;
  const int H = sl.H, W = sl.W;
  const int dlo = (k - 1) >> 1, elo = k >> 1;
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

constexpr int CC_TW = 32, CC_TH = 16;   // 512-pixel tiles, 1 thread per pixel

// Phase 1: tile-local union-find in shared memory, then publish slide-local ids of the local roots.
__global__ void __launch_bounds__(CC_TW* CC_TH)
    k_ccl_local(const uint8_t* __restrict__ mask_all, SegBatch sb, int conn8_fg, int* __restrict__ L_all) {
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

// Phase 2: merge across tile borders (and optionally tie border background to the frame node).
__global__ void __launch_bounds__(256)
    k_ccl_border(const uint8_t* __restrict__ mask_all, SegBatch sb, long long total, int conn8_fg, int frame,
                 int* __restrict__ L_all) {
  const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= total) return;
  const int s = find_slide(sb.px_base, sb.n, g);
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

// Phase 3: flatten + per-root area of class `cls` pixels (warp-aggregated atomics).
__global__ void __launch_bounds__(256)
    k_ccl_flatten_area(const uint8_t* __restrict__ mask_all, SegBatch sb, long long total, int cls,
                       int* __restrict__ L_all, int* __restrict__ area_all) {
  const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const bool live = g < total;
  int root = -1 - (int)(threadIdx.x & 31);
  bool act = false;
  int* area = nullptr;
  if (live) {
    const int s = find_slide(sb.px_base, sb.n, g);
    const SegSlide sl = sb.slides[s];
    const int idx = (int)(g - sb.px_base[s]);
    int* L = L_all + sl.lab_off;
    const int r = uf_find(L, idx);
    L[idx] = r;
    act = (mask_all + sl.mask_off)[idx] == cls;
    if (act) {
      // key must be unique across slides inside one warp: use the global label slot
      root = (int)(sl.lab_off + r);
      area = area_all;
    }
  }
  if (cls < 0) return;   // flatten only (warp-uniform)
  unsigned peers = __match_any_sync(0xffffffffu, root);
  if (act && (__ffs(peers) - 1) == (int)(threadIdx.x & 31)) atomicAdd(&area[root], __popc(peers));
}

// cls==1: remove foreground components with area < thr; cls==0: fill background components with area < thr
__global__ void __launch_bounds__(256)
    k_ccl_apply(uint8_t* __restrict__ mask_all, SegBatch sb, long long total, const int* __restrict__ L_all,
                const int* __restrict__ area_all, int cls) {
  const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= total) return;
  const int s = find_slide(sb.px_base, sb.n, g);
  const SegSlide sl = sb.slides[s];
  const int idx = (int)(g - sb.px_base[s]);
  uint8_t* mask = mask_all + sl.mask_off;
  const int thr = cls ? sl.min_obj : sl.min_hole;
  if (thr <= 0) return;
  // L may not be fully flattened for pixels processed by other blocks of the fused flatten/area pass
  // when this kernel runs (it is: separate launch), so one lookup suffices.
  if (mask[idx] == cls && area_all[sl.lab_off + L_all[sl.lab_off + idx]] < thr) mask[idx] = 1 - cls;
}

// outermost: a foreground component is outermost iff one of its pixels is on the image border or
// 4-adjacent to the background component tied to the frame node.  (L flattened, incl. frame node.)
__global__ void __launch_bounds__(256)
    k_outer_flag(const uint8_t* __restrict__ mask_all, SegBatch sb, long long total, const int* __restrict__ L_all,
                 int* __restrict__ flag_all) {
  const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= total) return;
  const int s = find_slide(sb.px_base, sb.n, g);
  const SegSlide sl = sb.slides[s];
  const int idx = (int)(g - sb.px_base[s]);
  const uint8_t* __restrict__ mask = mask_all + sl.mask_off;
  if (mask[idx] != 1) return;
  const int* __restrict__ L = L_all + sl.lab_off;
  const int H = sl.H, W = sl.W;
  const int gy = idx / W, gx = idx - gy * W;
  bool o = gx == 0 || gy == 0 || gx == W - 1 || gy == H - 1;
  if (!o) {
    const int fr = uf_find(L, H * W);
    o = (mask[idx - 1] == 0 && L[idx - 1] == fr) || (mask[idx + 1] == 0 && L[idx + 1] == fr) ||
        (mask[idx - W] == 0 && L[idx - W] == fr) || (mask[idx + W] == 0 && L[idx + W] == fr);
  }
  if (o) flag_all[sl.lab_off + L[idx]] = 1;
}
__global__ void __launch_bounds__(256)
    k_outer_apply(const uint8_t* __restrict__ mask_all, SegBatch sb, long long total, const int* __restrict__ L_all,
                  const int* __restrict__ flag_all, uint8_t* __restrict__ outer_all) {
  const long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= total) return;
  const int s = find_slide(sb.px_base, sb.n, g);
  const SegSlide sl = sb.slides[s];
  const int idx = (int)(g - sb.px_base[s]);
  const uint8_t m = (mask_all + sl.mask_off)[idx];
  (outer_all + sl.mask_off)[idx] = (m == 1 && flag_all[sl.lab_off + L_all[sl.lab_off + idx]]) ? 1 : 0;
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
  size_t off_desc, off_tile, off_ctile, off_px, off_tmp, off_L, off_aux, total;
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
  b.off_tmp = o;   o += align_up((size_t)b.npx, 256);
  b.off_L = o;     o += align_up(((size_t)b.npx + n) * sizeof(int), 256);
  b.off_aux = o;   o += align_up(((size_t)b.npx + n) * sizeof(int), 256);
  b.total = o;
  return b;
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
  // descriptor tables (host staging -> workspace head)
  std::vector<SegSlide> hs(n);
  std::vector<long long> tb(n + 1), cb(n + 1), pb(n + 1);
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
  WSI_CUDA_CHECK(cudaMemcpyAsync(ws + bl.off_desc, hs.data(), sizeof(SegSlide) * n, cudaMemcpyHostToDevice, stream));
  WSI_CUDA_CHECK(cudaMemcpyAsync(ws + bl.off_tile, tb.data(), sizeof(long long) * (n + 1), cudaMemcpyHostToDevice,
                                 stream));
  WSI_CUDA_CHECK(cudaMemcpyAsync(ws + bl.off_ctile, cb.data(), sizeof(long long) * (n + 1), cudaMemcpyHostToDevice,
                                 stream));
  WSI_CUDA_CHECK(cudaMemcpyAsync(ws + bl.off_px, pb.data(), sizeof(long long) * (n + 1), cudaMemcpyHostToDevice,
                                 stream));
  sb_out->slides = (const SegSlide*)(ws + bl.off_desc);
  sb_out->tile_base = (const long long*)(ws + bl.off_tile);
  sb_out->ctile_base = (const long long*)(ws + bl.off_ctile);
  sb_out->px_base = (const long long*)(ws + bl.off_px);
  sb_out->n = n;
  *bl_out = bl;
  return WSI_OK;
}

// 8-connected foreground labelling + frame adjacency -> outermost-component mask
static int seg_outermost(const SegBatch& sb, const BatchLayout& bl, const uint8_t* d_mask, uint8_t* d_outer, void* d_ws,
                         cudaStream_t stream) {
  uint8_t* ws = (uint8_t*)d_ws;
  int* L = (int*)(ws + bl.off_L);
  int* aux = (int*)(ws + bl.off_aux);
  const long long total = bl.npx;
  const size_t lab_bytes = ((size_t)bl.npx + sb.n) * sizeof(int);
  const int nb = wsi_cdiv(total, 256);
  k_ccl_local<<<(unsigned)bl.nctiles, CC_TW * CC_TH, 0, stream>>>(d_mask, sb, 1, L);
  k_ccl_border<<<nb, 256, 0, stream>>>(d_mask, sb, total, 1, 1, L);
  k_ccl_flatten_area<<<nb, 256, 0, stream>>>(d_mask, sb, total, -1, L, aux);
  WSI_CUDA_CHECK(cudaMemsetAsync(aux, 0, lab_bytes, stream));
  k_outer_flag<<<nb, 256, 0, stream>>>(d_mask, sb, total, L, aux);
  k_outer_apply<<<nb, 256, 0, stream>>>(d_mask, sb, total, L, aux, d_outer);
  wsi_count_launch(5);
  return WSI_OK;
}

int wsi_seg_tissue_batch_u8(const uint8_t* d_rgb, int n, const int* h_H, const int* h_W, const int* h_pitch,
                            const long long* h_rgb_off, const long long* h_mask_off, int median_filter_size,
                            int binary_threshold, int closing_kernel_size, const int* h_min_object_px,
                            const int* h_min_hole_px, uint8_t* d_mask, uint8_t* d_outer, void* d_ws, size_t ws_bytes,
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
  const long long total = bl.npx;
  const size_t lab_bytes = ((size_t)bl.npx + n) * sizeof(int);
  const int nb = wsi_cdiv(total, 256);
  int launches = 0;
  {
    int r = median_filter_size / 2;
    size_t smem = (size_t)(K1_TW + 2 * r) * (K1_TH + 2 * r);
    k_sat_median_thresh<<<(unsigned)bl.ntiles, K1_THREADS, smem, stream>>>(d_rgb, sb, median_filter_size,
                                                                           binary_threshold, tmp);
    ++launches;
  }
  {
    int k = closing_kernel_size, halo = k - 1;
    size_t smem = (size_t)(K1_TW + 2 * halo) * (K1_TH + 2 * halo) + (size_t)(K1_TW + halo) * (K1_TH + halo);
    k_closing<<<(unsigned)bl.ntiles, K1_THREADS, smem, stream>>>(tmp, sb, k, d_mask);
    ++launches;
  }
  // remove_small_objects (4-conn foreground, area < min) then remove_small_holes (4-conn background)
  for (int pass = 0; pass < 2; ++pass) {
    const int cls = pass == 0 ? 1 : 0;
    bool any = false;
    for (int s = 0; s < n; ++s) any = any || ((cls ? h_min_object_px[s] : h_min_hole_px[s]) > 0);
    if (!any) continue;
    k_ccl_local<<<(unsigned)bl.nctiles, CC_TW * CC_TH, 0, stream>>>(d_mask, sb, 0, L);
    k_ccl_border<<<nb, 256, 0, stream>>>(d_mask, sb, total, 0, 0, L);
    WSI_CUDA_CHECK(cudaMemsetAsync(aux, 0, lab_bytes, stream));
    k_ccl_flatten_area<<<nb, 256, 0, stream>>>(d_mask, sb, total, cls, L, aux);
    k_ccl_apply<<<nb, 256, 0, stream>>>(d_mask, sb, total, L, aux, cls);
    launches += 4;
  }
  wsi_count_launch(launches);
  if (d_outer) {
    rc = seg_outermost(sb, bl, d_mask, d_outer, d_ws, stream);
    if (rc) return rc;
  }
  WSI_CUDA_CHECK(cudaGetLastError());
  return WSI_OK;
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
  rc = seg_outermost(sb, bl, d_mask, d_outer, d_ws, stream);
  if (rc) return rc;
  WSI_CUDA_CHECK(cudaGetLastError());
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

}  // extern "C"
