// N3 / a13: per-tile nucleus post-processing of the cell-model path, on the device.
//
// Replaces (reference, huangch/wsinsight, wsinsight/modellib/tilefuse.py):
//   _proc_np_hv        :39-103   np >= 0.5, ndi.label + remove_small_objects, cv2.normalize(MINMAX) of the HV maps,
//                                cv2.Sobel(ksize = 21), 1 - normalize, ridge / marker, binary_fill_holes,
//                                MORPH_OPEN with the 5x5 ellipse, label + remove_small_objects, GaussianBlur(3x3),
//                                skimage.segmentation.watershed(dist, markers, mask)
//   _stitching_worker  :109-174  interior crop, per-instance bounding box, class-probability means, largest external
//                                contour (cv2.findContours RETR_EXTERNAL / CHAIN_APPROX_SIMPLE)
// One call processes one padded tile of finalize() (:280-415); nothing goes through the host.
//
// Arithmetic conventions (pinned against OpenCV on the CPU, oracle/cellpost.py + tests/test_oracle.py):
//  * cv2.normalize(NORM_MINMAX, dtype = CV_32F): scale = (double)(float)(1 / (max - min)) (0 when max - min <=
//    DBL_EPSILON), shift = -(double)(float)(min * scale); 32f source: fmaf(src, scale, shift); 64f source:
//    (float) fma(src, scale, shift).
//  * Sobel ksize 21: separable, getDerivKernels(…, 21) integer binomial kernels in double, BORDER_REFLECT_101; the
//    summation order of OpenCV's SIMD filters is not reproduced (differences ~1e-16 relative before the min-max
//    normalisation to float32).
//  * everything after `overall - (1.0 - blb)` is float64, as numpy promotes it.
//  * watershed: Meyer flooding pops pixels in order of their flood level L (L = image on markers, else
//    max(image, min over 4-neighbours of L)); a pixel takes the label of the neighbour with the smallest L.  Computed
//    as two monotone relaxations (L, then (hops, label) among the minimal-L neighbours), each one cooperative kernel
//    with grid-wide barriers.  Identical to the sequential flooding except on exact float ties between two floods.
#include "common.cuh"
#include "../../include/wsi_b200.h"
#include <cooperative_groups.h>
#include <cub/cub.cuh>
#include <math.h>
#include <vector>

namespace cg = cooperative_groups;

namespace {

constexpr int CP_MAXK = 16;

struct CPTile {
  const float* np; long long np_pitch;     // nuclei probability, elements per row
  const float* hv; long long hv_pitch;     // 2 interleaved floats per pixel
  const float* tp; long long tp_pitch;     // K interleaved floats per pixel
  int K, H, W;                             // padded tile
  int iy0, iy1, ix0, ix1;                  // interior slice inside the padded tile
  int gx0, gy0;                            // slide coordinates of the interior's first pixel
  int min_size;
};

// counters / flags living in device memory (ints)
enum {
  C_BIN_ANY = 0, C_NUM1, C_BLB_ANY, C_MARK_ANY, C_NUM2, C_NMARK, C_NINST, C_NPTS, C_ERR,
  C_FLAG0, C_FLAG1, C_FLAG2, C_HVMM0 /* 4 ordered-uint min/max of h, v */ = 12, C_INTS = 16
};
// doubles min/max of the two Sobel images as ordered uint64: [0] hmin [1] hmax [2] vmin [3] vmax
__device__ __forceinline__ unsigned ord_f32(float f) {
  unsigned u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float unord_f32(unsigned u) {
  return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}
__device__ __forceinline__ unsigned long long ord_f64(double d) {
  unsigned long long u = (unsigned long long)__double_as_longlong(d);
  return (u >> 63) ? ~u : (u | 0x8000000000000000ull);
}
__device__ __forceinline__ double unord_f64(unsigned long long u) {
  return __longlong_as_double((long long)((u >> 63) ? (u & 0x7fffffffffffffffull) : ~u));
}

__device__ __forceinline__ int reflect101(int i, int n) {
  if (n == 1) return 0;
  while (i < 0 || i >= n) i = i < 0 ? -i : 2 * n - 2 - i;
  return i;
}

// cv2.normalize(NORM_MINMAX, alpha = 0, beta = 1, dtype = CV_32F) coefficients
__device__ __forceinline__ void minmax_coeffs(double smin, double smax, double& scale, double& shift) {
  const double s = (smax - smin > 2.220446049250313e-16) ? 1.0 / (smax - smin) : 0.0;
  scale = (double)(float)s;
  shift = 0.0 - (double)(float)(smin * scale);
}

// ---- 1: blb_bin = np >= 0.5, min / max of the HV maps ----
__global__ void __launch_bounds__(256) k_cp_bin(CPTile t, uint8_t* __restrict__ bin, int* __restrict__ ctr) {
  const int n = t.H * t.W;
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  unsigned hmin = 0xffffffffu, hmax = 0u, vmin = 0xffffffffu, vmax = 0u;
  bool fg = false;
  if (p < n) {
    const int y = p / t.W, x = p - y * t.W;
    fg = t.np[(size_t)y * t.np_pitch + x] >= 0.5f;
    bin[p] = fg ? 1 : 0;
    const float2 hv = *reinterpret_cast<const float2*>(t.hv + (size_t)y * t.hv_pitch + 2 * (size_t)x);
    hmin = hmax = ord_f32(hv.x);
    vmin = vmax = ord_f32(hv.y);
  }
  for (int o = 16; o; o >>= 1) {
    hmin = min(hmin, __shfl_xor_sync(0xffffffffu, hmin, o));
    hmax = max(hmax, __shfl_xor_sync(0xffffffffu, hmax, o));
    vmin = min(vmin, __shfl_xor_sync(0xffffffffu, vmin, o));
    vmax = max(vmax, __shfl_xor_sync(0xffffffffu, vmax, o));
  }
  const unsigned any = __ballot_sync(0xffffffffu, fg);
  if ((threadIdx.x & 31) == 0) {
    unsigned* mm = reinterpret_cast<unsigned*>(ctr + C_HVMM0);
    atomicMin(&mm[0], hmin);
    atomicMax(&mm[1], hmax);
    atomicMin(&mm[2], vmin);
    atomicMax(&mm[3], vmax);
    if (any) ctr[C_BIN_ANY] = 1;
  }
}

// number of components: pixels of the class that are their own root
__global__ void __launch_bounds__(256)
    k_cp_count_roots(const uint8_t* __restrict__ mask, const int* __restrict__ L, int n, int* __restrict__ out) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  const bool r = p < n && mask[p] && L[p] == p;
  const unsigned b = __ballot_sync(0xffffffffu, r);
  if ((threadIdx.x & 31) == 0 && b) atomicAdd(out, __popc(b));
}

// ---- 2: blb = components of blb_bin that survive remove_small_objects (applied only when there are > 1) ----
__global__ void __launch_bounds__(256)
    k_cp_blb(const uint8_t* __restrict__ bin, const int* __restrict__ L1, const int* __restrict__ area1, int n, int min_size,
             uint8_t* __restrict__ blb, int* __restrict__ ctr) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  bool v = false;
  if (p < n) {
    v = bin[p] && !(ctr[C_NUM1] > 1 && area1[L1[p]] < min_size);
    blb[p] = v ? 1 : 0;
  }
  if (__ballot_sync(0xffffffffu, v) && (threadIdx.x & 31) == 0) ctr[C_BLB_ANY] = 1;
}

// ---- 3: Sobel ksize 21 (separable, float64) of the min-max normalised HV maps ----
__constant__ double c_kd[21];   // derivative kernel   (getDerivKernels order 1)
__constant__ double c_ks[21];   // smoothing kernel    (order 0)

// row pass of both images: tmpH = deriv_x(h_dir), tmpV = smooth_x(v_dir)
__global__ void __launch_bounds__(256)
    k_cp_sobel_row(CPTile t, const int* __restrict__ ctr, double* __restrict__ tmpH, double* __restrict__ tmpV) {
  const int n = t.H * t.W;
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n || !ctr[C_BLB_ANY]) return;
  const unsigned* mm = reinterpret_cast<const unsigned*>(ctr + C_HVMM0);
  double sc, sh;
  minmax_coeffs((double)unord_f32(mm[0]), (double)unord_f32(mm[1]), sc, sh);
  const float hsc = (float)sc, hsh = (float)sh;
  minmax_coeffs((double)unord_f32(mm[2]), (double)unord_f32(mm[3]), sc, sh);
  const float vsc = (float)sc, vsh = (float)sh;
  const int y = p / t.W, x = p - y * t.W;
  const float* row = t.hv + (size_t)y * t.hv_pitch;
  // OpenCV's separable filters evaluate symmetric / anti-symmetric kernels in the folded form
  // k[0] * S[0] + sum_k k[k] * (S[+k] +- S[-k]): a derivative along an axis the image does not vary on is then exactly 0
  auto hv_at = [&](int xx) { return *reinterpret_cast<const float2*>(row + 2 * (size_t)reflect101(xx, t.W)); };
  const float2 c0 = hv_at(x);
  double sH = 0.0, sV = c_ks[10] * (double)fmaf(c0.y, vsc, vsh);
#pragma unroll
  for (int k = 1; k <= 10; ++k) {
    const float2 a = hv_at(x + k), b = hv_at(x - k);
    sH += c_kd[10 + k] * ((double)fmaf(a.x, hsc, hsh) - (double)fmaf(b.x, hsc, hsh));
    sV += c_ks[10 + k] * ((double)fmaf(a.y, vsc, vsh) + (double)fmaf(b.y, vsc, vsh));
  }
  tmpH[p] = sH;
  tmpV[p] = sV;
}

// column pass: sobelh = smooth_y(tmpH), sobelv = deriv_y(tmpV); min / max of both
__global__ void __launch_bounds__(256)
    k_cp_sobel_col(int H, int W, const int* __restrict__ ctr, const double* __restrict__ tmpH,
                   const double* __restrict__ tmpV, double* __restrict__ sobH, double* __restrict__ sobV,
                   unsigned long long* __restrict__ mm64) {
  const int n = H * W;
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (!ctr[C_BLB_ANY]) return;
  unsigned long long a0 = ~0ull, a1 = 0ull, b0 = ~0ull, b1 = 0ull;
  if (p < n) {
    const int y = p / W, x = p - y * W;
    double sH = c_ks[10] * tmpH[(size_t)y * W + x], sV = 0.0;
#pragma unroll
    for (int k = 1; k <= 10; ++k) {
      const size_t ia = (size_t)reflect101(y + k, H) * W + x, ib = (size_t)reflect101(y - k, H) * W + x;
      sH += c_ks[10 + k] * (tmpH[ia] + tmpH[ib]);
      sV += c_kd[10 + k] * (tmpV[ia] - tmpV[ib]);
    }
    sobH[p] = sH;
    sobV[p] = sV;
    a0 = a1 = ord_f64(sH);
    b0 = b1 = ord_f64(sV);
  }
  for (int o = 16; o; o >>= 1) {
    a0 = min(a0, __shfl_xor_sync(0xffffffffu, a0, o));
    a1 = max(a1, __shfl_xor_sync(0xffffffffu, a1, o));
    b0 = min(b0, __shfl_xor_sync(0xffffffffu, b0, o));
    b1 = max(b1, __shfl_xor_sync(0xffffffffu, b1, o));
  }
  if ((threadIdx.x & 31) == 0) {
    atomicMin(&mm64[0], a0);
    atomicMax(&mm64[1], a1);
    atomicMin(&mm64[2], b0);
    atomicMax(&mm64[3], b1);
  }
}

// ---- 4: overall, dist0 = (1 - overall) * blb, marker0 = blb & !(overall >= 0.4) ----
__global__ void __launch_bounds__(256)
    k_cp_overall(int n, const double* __restrict__ sobH, const double* __restrict__ sobV,
                 const unsigned long long* __restrict__ mm64, const uint8_t* __restrict__ blb, double* __restrict__ dist0,
                 uint8_t* __restrict__ marker0, int* __restrict__ ctr) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (!ctr[C_BLB_ANY]) {            // no nucleus in the tile: nothing was filtered, the marker image is empty
    if (p < n) marker0[p] = 0;
    return;
  }
  bool mk = false;
  if (p < n) {
    double sc, sh;
    minmax_coeffs(unord_f64(mm64[0]), unord_f64(mm64[1]), sc, sh);
    const float nh = 1.0f - (float)fma(sobH[p], sc, sh);
    minmax_coeffs(unord_f64(mm64[2]), unord_f64(mm64[3]), sc, sh);
    const float nv = 1.0f - (float)fma(sobV[p], sc, sh);
    const double b = (double)blb[p];
    double ov = (double)fmaxf(nh, nv) - (1.0 - b);
    if (ov < 0.0) ov = 0.0;
    dist0[p] = (1.0 - ov) * b;
    mk = blb[p] && !(ov >= 0.4);
    marker0[p] = mk ? 1 : 0;
  }
  if (__ballot_sync(0xffffffffu, mk) && (threadIdx.x & 31) == 0) ctr[C_MARK_ANY] = 1;
}

// ---- 5: dist = -GaussianBlur(dist0, (3, 3), 0): kernel [0.25, 0.5, 0.25], BORDER_REFLECT_101, float64 ----
__global__ void __launch_bounds__(256)
    k_cp_gauss(int H, int W, const int* __restrict__ ctr, const double* __restrict__ src, double* __restrict__ dst) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= H * W || !ctr[C_BLB_ANY]) return;
  const int y = p / W, x = p - y * W;
  const int xl = reflect101(x - 1, W), xr = reflect101(x + 1, W);
  double r[3];
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    const double* row = src + (size_t)reflect101(y + j - 1, H) * W;
    r[j] = (row[xl] + row[xr]) * 0.25 + row[x] * 0.5;
  }
  dst[p] = -((r[0] + r[2]) * 0.25 + r[1] * 0.5);
}

// ---- 6: binary_fill_holes: background not connected to the frame becomes marker ----
__global__ void __launch_bounds__(256)
    k_cp_fill(int n, const uint8_t* __restrict__ marker0, const int* __restrict__ Lf, uint8_t* __restrict__ filled) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n) return;
  int fr = Lf[n];
  while (Lf[fr] != fr) fr = Lf[fr];
  filled[p] = (marker0[p] || Lf[p] != fr) ? 1 : 0;
}

// ---- 7: erosion / dilation with cv2.getStructuringElement(MORPH_ELLIPSE, (5, 5)); erosion ignores the outside ----
__global__ void __launch_bounds__(256)
    k_cp_morph5(int H, int W, const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, int erode) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= H * W) return;
  const int y = p / W, x = p - y * W;
  bool v = erode != 0;
#pragma unroll
  for (int dy = -2; dy <= 2; ++dy) {
    const int hw = (dy == -2 || dy == 2) ? 0 : 2;       // rows 0 and 4 of the ellipse hold the centre column only
    const int yy = y + dy;
    for (int dx = -hw; dx <= hw; ++dx) {
      const int xx = x + dx;
      const bool in = yy >= 0 && yy < H && xx >= 0 && xx < W;
      if (erode) v = v && (!in || src[(size_t)yy * W + xx]);
      else v = v || (in && src[(size_t)yy * W + xx]);
    }
  }
  dst[p] = v ? 1 : 0;
}

// ---- 8: marker label roots: opened components that survive remove_small_objects, or (no marker at all) label(blb) ----
__global__ void __launch_bounds__(256)
    k_cp_marker_roots(int n, const int* __restrict__ ctr, const uint8_t* __restrict__ opened, const int* __restrict__ L2,
                      const int* __restrict__ area2, const uint8_t* __restrict__ blb, const int* __restrict__ L1, int min_size,
                      int* __restrict__ root_of, int* __restrict__ is_root) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n) return;
  int r = -1;
  if (ctr[C_BLB_ANY]) {
    if (ctr[C_MARK_ANY]) {
      if (opened[p] && !(ctr[C_NUM2] > 1 && area2[L2[p]] < min_size)) r = L2[p];
    } else if (blb[p]) {
      r = L1[p];
    }
  }
  root_of[p] = r;
  is_root[p] = (r == p) ? 1 : 0;
}

// ---- 9: watershed ----
constexpr unsigned long long KEY_NONE = ~0ull;

__global__ void __launch_bounds__(256)
    k_cp_ws_init(int n, const int* __restrict__ root_of, const int* __restrict__ rank, const double* __restrict__ dist,
                 double* __restrict__ Lv, unsigned long long* __restrict__ key, int* __restrict__ ctr) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n) return;
  const int r = root_of[p];
  if (r >= 0) {
    const int id = rank[r] + 1;          // dense ids in raster order of the component's first pixel (== scipy's numbering)
    Lv[p] = dist[p];
    key[p] = (unsigned long long)(unsigned)id;
    if (r == p) atomicMax(&ctr[C_NMARK], id);
  } else {
    Lv[p] = INFINITY;
    key[p] = KEY_NONE;
  }
}

// Phase 1: flood levels.  In-place monotone relaxation; a 3-slot rotating "changed" flag ends the loop.
__global__ void __launch_bounds__(256)
    k_cp_ws_levels(int H, int W, const uint8_t* __restrict__ blb, const int* __restrict__ root_of,
                   const double* __restrict__ dist, double* Lv, int* ctr) {
  cg::grid_group grid = cg::this_grid();
  const int n = H * W;
  volatile double* L = Lv;
  volatile int* flag = ctr + C_FLAG0;
  for (int it = 0;; ++it) {
    if (grid.thread_rank() == 0) flag[(it + 1) % 3] = 0;
    bool ch = false;
    for (int p = (int)grid.thread_rank(); p < n; p += (int)grid.size()) {
      if (!blb[p] || root_of[p] >= 0) continue;
      const int y = p / W, x = p - y * W;
      double m = INFINITY;
      if (y > 0) m = fmin(m, L[p - W]);
      if (x > 0) m = fmin(m, L[p - 1]);
      if (x < W - 1) m = fmin(m, L[p + 1]);
      if (y < H - 1) m = fmin(m, L[p + W]);
      const double c = fmax(dist[p], m);
      if (c < L[p]) {
        L[p] = c;
        ch = true;
      }
    }
    if (ch) flag[it % 3] = 1;
    __threadfence();
    grid.sync();
    if (!flag[it % 3]) break;
  }
}

// Phase 2: (hops << 32 | label), propagated only from the neighbours whose level is the minimum around the pixel.
__global__ void __launch_bounds__(256)
    k_cp_ws_labels(int H, int W, const uint8_t* __restrict__ blb, const int* __restrict__ root_of,
                   const double* __restrict__ Lv, unsigned long long* key_, int* ctr) {
  cg::grid_group grid = cg::this_grid();
  const int n = H * W;
  volatile unsigned long long* key = key_;
  volatile int* flag = ctr + C_FLAG0;
  for (int it = 0;; ++it) {
    if (grid.thread_rank() == 0) flag[(it + 1) % 3] = 0;
    bool ch = false;
    for (int p = (int)grid.thread_rank(); p < n; p += (int)grid.size()) {
      if (!blb[p] || root_of[p] >= 0) continue;
      const int y = p / W, x = p - y * W;
      const double l0 = y > 0 ? Lv[p - W] : INFINITY, l1 = x > 0 ? Lv[p - 1] : INFINITY;
      const double l2 = x < W - 1 ? Lv[p + 1] : INFINITY, l3 = y < H - 1 ? Lv[p + W] : INFINITY;
      const double m = fmin(fmin(l0, l1), fmin(l2, l3));
      if (!(m < INFINITY)) continue;
      unsigned long long best = key[p];
      const unsigned long long one = 1ull << 32;
      unsigned long long k;
      if (l0 == m && (k = key[p - W]) != KEY_NONE && k + one < best) best = k + one;
      if (l1 == m && (k = key[p - 1]) != KEY_NONE && k + one < best) best = k + one;
      if (l2 == m && (k = key[p + 1]) != KEY_NONE && k + one < best) best = k + one;
      if (l3 == m && (k = key[p + W]) != KEY_NONE && k + one < best) best = k + one;
      if (best < key[p]) {
        key[p] = best;
        ch = true;
      }
    }
    if (ch) flag[it % 3] = 1;
    __threadfence();
    grid.sync();
    if (!flag[it % 3]) break;
  }
}

// ---- 10: instance map + per-instance statistics over the interior ----
struct CPStats {
  int* cnt;        // [nmark + 1]
  int* bbox;       // [nmark + 1][4] minx, miny, maxx, maxy (interior-local)
  int* first;      // [nmark + 1] first raster pixel (interior-local linear index)
  double* cls;     // [nmark + 1][K]
};

__global__ void __launch_bounds__(256)
    k_cp_stats_init(int m, int K, CPStats s) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i > m) return;
  s.cnt[i] = 0;
  s.bbox[4 * i] = 0x7fffffff; s.bbox[4 * i + 1] = 0x7fffffff; s.bbox[4 * i + 2] = -1; s.bbox[4 * i + 3] = -1;
  s.first[i] = 0x7fffffff;
  for (int k = 0; k < K; ++k) s.cls[(size_t)i * K + k] = 0.0;
}

__global__ void __launch_bounds__(256)
    k_cp_inst(CPTile t, const unsigned long long* __restrict__ key, int* __restrict__ inst, CPStats s, int max_marks) {
  const int n = t.H * t.W;
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n) return;
  const unsigned long long k = key[p];
  const int id = k == KEY_NONE ? 0 : (int)(unsigned)(k & 0xffffffffull);
  inst[p] = id;
  if (id <= 0 || id > max_marks) return;
  const int y = p / t.W, x = p - y * t.W;
  if (y < t.iy0 || y >= t.iy1 || x < t.ix0 || x >= t.ix1) return;
  const int ly = y - t.iy0, lx = x - t.ix0, iw = t.ix1 - t.ix0;
  atomicAdd(&s.cnt[id], 1);
  atomicMin(&s.bbox[4 * id], lx);
  atomicMin(&s.bbox[4 * id + 1], ly);
  atomicMax(&s.bbox[4 * id + 2], lx);
  atomicMax(&s.bbox[4 * id + 3], ly);
  atomicMin(&s.first[id], ly * iw + lx);
  const float* tp = t.tp + (size_t)y * t.tp_pitch + (size_t)x * t.K;
  for (int c = 0; c < t.K; ++c) atomicAdd(&s.cls[(size_t)id * t.K + c], (double)tp[c]);
}

// instances that own interior pixels, in id order: out index = exclusive scan of (cnt > 0)
__global__ void __launch_bounds__(256) k_cp_valid(int m, const int* __restrict__ cnt, int* __restrict__ valid) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i <= m) valid[i] = (i > 0 && cnt[i] > 0) ? 1 : 0;
}

// OpenCV border following (contours.cpp icvFetchContour, outer border, CHAIN_APPROX_SIMPLE) of the 8-connected
// component of instance `id` that holds its first raster pixel, inside the interior crop.  emit == nullptr: count only.
__device__ int cp_trace(const int* __restrict__ inst, const CPTile& t, int id, int first, int2* emit, int cap, int ox, int oy) {
  const int iw = t.ix1 - t.ix0, ih = t.iy1 - t.iy0;
  const int dxs[8] = {1, 1, 0, -1, -1, -1, 0, 1}, dys[8] = {0, -1, -1, -1, 0, 1, 1, 1};
  auto at = [&](int x, int y) {
    return x >= 0 && y >= 0 && x < iw && y < ih && inst[(size_t)(y + t.iy0) * t.W + (x + t.ix0)] == id;
  };
  const int y0 = first / iw, x0 = first - y0 * iw;
  int npts = 0;
  auto put = [&](int x, int y) {
    if (emit && npts < cap) emit[npts] = make_int2(x + ox, y + oy);
    ++npts;
  };
  int s = 4;
  const int s_stop = 4;
  do {
    s = (s - 1) & 7;
  } while (!at(x0 + dxs[s], y0 + dys[s]) && s != s_stop);
  if (s == s_stop) {
    put(x0, y0);
    return npts;
  }
  const int x1 = x0 + dxs[s], y1 = y0 + dys[s];
  int x3 = x0, y3 = y0, prev_s = s ^ 4, px = x0, py = y0;
  for (int guard = 0; guard < (1 << 24); ++guard) {
    int x4, y4;
    do {
      s = (s + 1) & 7;
      x4 = x3 + dxs[s];
      y4 = y3 + dys[s];
    } while (!at(x4, y4));
    if (s != prev_s) put(px, py);
    prev_s = s;
    px += dxs[s];
    py += dys[s];
    if (x4 == x0 && y4 == y0 && x3 == x1 && y3 == y1) break;
    x3 = x4;
    y3 = y4;
    s = (s + 4) & 7;
  }
  return npts;
}

// pass A: per valid instance write bbox / class means and count its contour points
__global__ void __launch_bounds__(128)
    k_cp_emit(CPTile t, int m, const int* __restrict__ valid, const int* __restrict__ vrank, CPStats s,
              const int* __restrict__ inst, int max_inst, int* __restrict__ out_bbox, float* __restrict__ out_cls,
              int* __restrict__ out_npts, int* __restrict__ ctr) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i > m || i == 0 || !valid[i]) return;
  const int o = vrank[i];
  atomicMax(&ctr[C_NINST], o + 1);
  if (o >= max_inst) {
    ctr[C_ERR] = 1;
    return;
  }
  const int* b = s.bbox + 4 * i;
  out_bbox[4 * o] = b[0] + t.gx0;
  out_bbox[4 * o + 1] = b[1] + t.gy0;
  out_bbox[4 * o + 2] = b[2] - b[0] + 1;
  out_bbox[4 * o + 3] = b[3] - b[1] + 1;
  const double den = (double)s.cnt[i];
  for (int c = 0; c < t.K; ++c) out_cls[(size_t)o * t.K + c] = (float)(s.cls[(size_t)i * t.K + c] / den);
  const int np = cp_trace(inst, t, i, s.first[i], nullptr, 0, 0, 0);
  out_npts[o] = np >= 3 ? np : 0;        // polygons with < 3 vertices are dropped (tilefuse.py:166-167)
}

// pass B: write the contour points at their scanned offsets
__global__ void __launch_bounds__(128)
    k_cp_poly(CPTile t, int m, const int* __restrict__ valid, const int* __restrict__ vrank, CPStats s,
              const int* __restrict__ inst, int max_inst, const int* __restrict__ out_npts, const int* __restrict__ poly_off,
              int max_pts, int2* __restrict__ poly_xy, int* __restrict__ ctr) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i > m || i == 0 || !valid[i]) return;
  const int o = vrank[i];
  if (o >= max_inst) return;
  const int np = out_npts[o], off = poly_off[o];
  if (o == ctr[C_NINST] - 1) ctr[C_NPTS] = off + np;
  if (np == 0) return;
  if (off + np > max_pts) {
    ctr[C_ERR] = 2;
    return;
  }
  cp_trace(inst, t, i, s.first[i], poly_xy + off, np, t.gx0, t.gy0);
}

}  // namespace

// forward declaration of the labelling entry in seg.cu
extern "C" int wsi_ccl_label_u8(const uint8_t* d_mask, int H, int W, int conn8_fg, int frame, int area_cls, int** d_L_out,
                                int** d_area_out, void* d_ws, size_t ws_bytes, void* stream);
extern "C" size_t wsi_seg_workspace_bytes(int H, int W);

namespace {

__global__ void k_cp_ctr_init(int* ctr, unsigned long long* mm64) {
  const int i = threadIdx.x;
  if (i < C_INTS) ctr[i] = 0;
  __syncthreads();
  if (i == 0) {
    unsigned* mm = reinterpret_cast<unsigned*>(ctr + C_HVMM0);
    mm[0] = 0xffffffffu; mm[1] = 0u; mm[2] = 0xffffffffu; mm[3] = 0u;
    mm64[0] = ~0ull; mm64[1] = 0ull; mm64[2] = ~0ull; mm64[3] = 0ull;
  }
}

size_t up256(size_t v) { return (v + 255) / 256 * 256; }

struct CPLayout {
  size_t ctr, mm64, u8[6], f64[4], key, i32[4], ccl[3], stats_cnt, stats_bbox, stats_first, stats_cls, valid, vrank, npts_off,
      cub, total;
  size_t ccl_bytes, cub_bytes;
  int max_marks;
};

CPLayout cp_layout(int H, int W, int K, int max_inst) {
  CPLayout l;
  const size_t n = (size_t)H * W;
  size_t o = 0;
  auto take = [&](size_t bytes) {
    size_t r = o;
    o += up256(bytes);
    return r;
  };
  l.max_marks = (int)(n / 2 + 16);      // a binary image has at most ceil(n / 2) 4-connected components
  l.ctr = take(sizeof(int) * C_INTS);
  l.mm64 = take(sizeof(unsigned long long) * 4);
  for (int i = 0; i < 6; ++i) l.u8[i] = take(n);
  for (int i = 0; i < 4; ++i) l.f64[i] = take(n * 8);
  l.key = take(n * 8);
  for (int i = 0; i < 4; ++i) l.i32[i] = take(n * 4);
  l.ccl_bytes = wsi_seg_workspace_bytes(H, W);
  for (int i = 0; i < 3; ++i) l.ccl[i] = take(l.ccl_bytes);
  const size_t m1 = (size_t)l.max_marks + 1;
  l.stats_cnt = take(m1 * 4);
  l.stats_bbox = take(m1 * 16);
  l.stats_first = take(m1 * 4);
  l.stats_cls = take(m1 * 8 * (size_t)K);
  l.valid = take(m1 * 4);
  l.vrank = take(m1 * 4);
  l.npts_off = take(((size_t)max_inst + 1) * 4);
  size_t c1 = 0, c2 = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, c1, (const int*)nullptr, (int*)nullptr, (int)n);
  cub::DeviceScan::ExclusiveSum(nullptr, c2, (const int*)nullptr, (int*)nullptr, (int)m1);
  l.cub_bytes = (c1 > c2 ? c1 : c2) + 256;
  l.cub = take(l.cub_bytes);
  l.total = o;
  return l;
}

bool g_sobel_ready = false;

int upload_sobel_kernels() {
  if (g_sobel_ready) return WSI_OK;
  // cv::getSobelKernels(ksize = 21): smoothing = C(20, i), first derivative = C(19, i - 1) - C(19, i)
  double c19[22], ks[21], kd[21];
  c19[0] = 1.0;
  for (int i = 1; i <= 19; ++i) c19[i] = c19[i - 1] * (double)(19 - i + 1) / (double)i;
  double c20 = 1.0;
  for (int i = 0; i <= 20; ++i) {
    ks[i] = __builtin_rint(c20);
    c20 = c20 * (double)(20 - i) / (double)(i + 1);
    const double a = i >= 1 ? __builtin_rint(c19[i - 1]) : 0.0, b = i <= 19 ? __builtin_rint(c19[i]) : 0.0;
    kd[i] = a - b;
  }
  WSI_CUDA_CHECK(cudaMemcpyToSymbol(c_kd, kd, sizeof(kd)));
  WSI_CUDA_CHECK(cudaMemcpyToSymbol(c_ks, ks, sizeof(ks)));
  g_sobel_ready = true;
  return WSI_OK;
}

int coop_blocks(const void* fn, int threads) {
  int dev = 0, sms = 0, per = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per, fn, threads, 0);
  if (per < 1) per = 1;
  return sms * per;
}

}  // namespace

extern "C" {

size_t wsi_cellpost_workspace_bytes(int H, int W, int K, int max_instances) {
  if (H <= 0 || W <= 0 || K <= 0 || max_instances < 0) return 0;
  return cp_layout(H, W, K, max_instances).total;
}

int wsi_cellpost_tile(const float* d_np, long long np_pitch, const float* d_hv, long long hv_pitch, const float* d_tp,
                      long long tp_pitch, int K, int H, int W, int iy0, int iy1, int ix0, int ix1, int gx0, int gy0,
                      int min_object_size, int max_instances, int max_poly_points, int* d_counts, int* d_bbox,
                      float* d_cls_mean, int* d_poly_npts, int* d_poly_off, int* d_poly_xy, int* d_inst_map, void* d_ws,
                      size_t ws_bytes, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  WSI_ARG_CHECK(d_np && d_hv && d_tp && d_counts && d_bbox && d_cls_mean && d_poly_npts && d_poly_off && d_poly_xy && d_ws,
                "wsi_cellpost_tile: null pointer");
  WSI_ARG_CHECK(H >= 1 && W >= 1 && (long long)H * W < (1ll << 30) && K >= 1 && K <= CP_MAXK,
                "wsi_cellpost_tile: bad sizes (K <= %d)", CP_MAXK);
  WSI_ARG_CHECK(0 <= iy0 && iy0 <= iy1 && iy1 <= H && 0 <= ix0 && ix0 <= ix1 && ix1 <= W,
                "wsi_cellpost_tile: the interior slice must lie inside the padded tile");
  WSI_ARG_CHECK(max_instances >= 1 && max_poly_points >= 1 && min_object_size >= 0, "wsi_cellpost_tile: bad capacities");
  WSI_ARG_CHECK((((uintptr_t)d_hv) & 7) == 0 && (hv_pitch & 1) == 0, "wsi_cellpost_tile: the hv map must be 8-byte aligned");
  const CPLayout l = cp_layout(H, W, K, max_instances);
  WSI_ARG_CHECK(ws_bytes >= l.total, "wsi_cellpost_tile: workspace too small (%zu < %zu)", ws_bytes, l.total);
  int rc = upload_sobel_kernels();
  if (rc) return rc;
  uint8_t* ws = (uint8_t*)d_ws;
  int* ctr = (int*)(ws + l.ctr);
  unsigned long long* mm64 = (unsigned long long*)(ws + l.mm64);
  uint8_t *bin = ws + l.u8[0], *blb = ws + l.u8[1], *marker0 = ws + l.u8[2], *filled = ws + l.u8[3], *eroded = ws + l.u8[4],
          *opened = ws + l.u8[5];
  double *tmpH = (double*)(ws + l.f64[0]), *tmpV = (double*)(ws + l.f64[1]), *sobH = (double*)(ws + l.f64[2]),
         *sobV = (double*)(ws + l.f64[3]);
  double *dist0 = tmpH, *dist = tmpV, *Lv = sobH;     // reused once the Sobel images are consumed
  unsigned long long* key = (unsigned long long*)(ws + l.key);
  int *root_of = (int*)(ws + l.i32[0]), *is_root = (int*)(ws + l.i32[1]), *rank = (int*)(ws + l.i32[2]);
  int* inst = d_inst_map ? d_inst_map : (int*)(ws + l.i32[3]);
  CPStats st;
  st.cnt = (int*)(ws + l.stats_cnt);
  st.bbox = (int*)(ws + l.stats_bbox);
  st.first = (int*)(ws + l.stats_first);
  st.cls = (double*)(ws + l.stats_cls);
  int *valid = (int*)(ws + l.valid), *vrank = (int*)(ws + l.vrank);
  CPTile t;
  t.np = d_np; t.np_pitch = np_pitch; t.hv = d_hv; t.hv_pitch = hv_pitch; t.tp = d_tp; t.tp_pitch = tp_pitch;
  t.K = K; t.H = H; t.W = W; t.iy0 = iy0; t.iy1 = iy1; t.ix0 = ix0; t.ix1 = ix1; t.gx0 = gx0; t.gy0 = gy0;
  t.min_size = min_object_size;
  const int n = H * W;
  const int nb = wsi_cdiv(n, 256);
  int launches = 0;
  k_cp_ctr_init<<<1, 32, 0, stream>>>(ctr, mm64);
  k_cp_bin<<<nb, 256, 0, stream>>>(t, bin, ctr);
  int *L1 = nullptr, *area1 = nullptr, *Lf = nullptr, *L2 = nullptr, *area2 = nullptr;
  rc = wsi_ccl_label_u8(bin, H, W, 0, 0, 1, &L1, &area1, ws + l.ccl[0], l.ccl_bytes, stream);
  if (rc) return rc;
  k_cp_count_roots<<<nb, 256, 0, stream>>>(bin, L1, n, ctr + C_NUM1);
  k_cp_blb<<<nb, 256, 0, stream>>>(bin, L1, area1, n, min_object_size, blb, ctr);
  k_cp_sobel_row<<<nb, 256, 0, stream>>>(t, ctr, tmpH, tmpV);
  k_cp_sobel_col<<<nb, 256, 0, stream>>>(H, W, ctr, tmpH, tmpV, sobH, sobV, mm64);
  k_cp_overall<<<nb, 256, 0, stream>>>(n, sobH, sobV, mm64, blb, dist0, marker0, ctr);
  k_cp_gauss<<<nb, 256, 0, stream>>>(H, W, ctr, dist0, dist);
  launches += 8;
  rc = wsi_ccl_label_u8(marker0, H, W, 0, 1, -1, &Lf, nullptr, ws + l.ccl[1], l.ccl_bytes, stream);
  if (rc) return rc;
  k_cp_fill<<<nb, 256, 0, stream>>>(n, marker0, Lf, filled);
  k_cp_morph5<<<nb, 256, 0, stream>>>(H, W, filled, eroded, 1);
  k_cp_morph5<<<nb, 256, 0, stream>>>(H, W, eroded, opened, 0);
  rc = wsi_ccl_label_u8(opened, H, W, 0, 0, 1, &L2, &area2, ws + l.ccl[2], l.ccl_bytes, stream);
  if (rc) return rc;
  k_cp_count_roots<<<nb, 256, 0, stream>>>(opened, L2, n, ctr + C_NUM2);
  k_cp_marker_roots<<<nb, 256, 0, stream>>>(n, ctr, opened, L2, area2, blb, L1, min_object_size, root_of, is_root);
  size_t cub_bytes = l.cub_bytes;
  WSI_CUDA_CHECK(cub::DeviceScan::ExclusiveSum(ws + l.cub, cub_bytes, is_root, rank, n, stream));
  k_cp_ws_init<<<nb, 256, 0, stream>>>(n, root_of, rank, dist, Lv, key, ctr);
  launches += 7;
  {
    // cooperative kernels: every CTA resident, grid-wide barrier between relaxation sweeps
    const int g1 = coop_blocks((const void*)k_cp_ws_levels, 256), g2 = coop_blocks((const void*)k_cp_ws_labels, 256);
    const uint8_t* blb_c = blb;
    const int* root_c = root_of;
    const double* dist_c = dist;
    const double* Lv_c = Lv;
    int Hh = H, Ww = W;
    void* a1[] = {&Hh, &Ww, &blb_c, &root_c, &dist_c, &Lv, &ctr};
    WSI_CUDA_CHECK(cudaMemsetAsync(ctr + C_FLAG0, 0, 3 * sizeof(int), stream));
    WSI_CUDA_CHECK(cudaLaunchCooperativeKernel((const void*)k_cp_ws_levels, dim3(g1), dim3(256), a1, 0, stream));
    void* a2[] = {&Hh, &Ww, &blb_c, &root_c, &Lv_c, &key, &ctr};
    WSI_CUDA_CHECK(cudaMemsetAsync(ctr + C_FLAG0, 0, 3 * sizeof(int), stream));
    WSI_CUDA_CHECK(cudaLaunchCooperativeKernel((const void*)k_cp_ws_labels, dim3(g2), dim3(256), a2, 0, stream));
    launches += 2;
  }
  const int m = l.max_marks;
  k_cp_stats_init<<<wsi_cdiv(m + 1, 256), 256, 0, stream>>>(m, K, st);
  k_cp_inst<<<nb, 256, 0, stream>>>(t, key, inst, st, m);
  k_cp_valid<<<wsi_cdiv(m + 1, 256), 256, 0, stream>>>(m, st.cnt, valid);
  cub_bytes = l.cub_bytes;
  WSI_CUDA_CHECK(cub::DeviceScan::ExclusiveSum(ws + l.cub, cub_bytes, valid, vrank, m + 1, stream));
  WSI_CUDA_CHECK(cudaMemsetAsync(d_poly_npts, 0, sizeof(int) * (size_t)max_instances, stream));
  k_cp_emit<<<wsi_cdiv(m + 1, 128), 128, 0, stream>>>(t, m, valid, vrank, st, inst, max_instances, d_bbox, d_cls_mean,
                                                        d_poly_npts, ctr);
  cub_bytes = l.cub_bytes;
  WSI_CUDA_CHECK(cub::DeviceScan::ExclusiveSum(ws + l.cub, cub_bytes, d_poly_npts, d_poly_off, max_instances, stream));
  k_cp_poly<<<wsi_cdiv(m + 1, 128), 128, 0, stream>>>(t, m, valid, vrank, st, inst, max_instances, d_poly_npts, d_poly_off,
                                                        max_poly_points, reinterpret_cast<int2*>(d_poly_xy), ctr);
  launches += 8;
  // counts for the caller: {instances, polygon points, error flag, markers}
  WSI_CUDA_CHECK(cudaMemcpyAsync(d_counts, ctr + C_NINST, sizeof(int), cudaMemcpyDeviceToDevice, stream));
  WSI_CUDA_CHECK(cudaMemcpyAsync(d_counts + 1, ctr + C_NPTS, sizeof(int), cudaMemcpyDeviceToDevice, stream));
  WSI_CUDA_CHECK(cudaMemcpyAsync(d_counts + 2, ctr + C_ERR, sizeof(int), cudaMemcpyDeviceToDevice, stream));
  WSI_CUDA_CHECK(cudaMemcpyAsync(d_counts + 3, ctr + C_NMARK, sizeof(int), cudaMemcpyDeviceToDevice, stream));
  wsi_count_launch(launches);
  WSI_CUDA_CHECK(cudaGetLastError());
  return WSI_OK;
}

}  // extern "C"
