// K3: fused patch data path  (crop RGBA->RGB, PIL-exact antialiased bilinear resize,
// ToTensor, Normalize) from a device-resident slide into bf16 NHWC4 batches.
//
// Replaces (reference, huangch/wsinsight):
//   wsinsight/modellib/data.py:270-281        read_region(...).convert("RGB")  (zero padded OOB)
//   wsinsight/modellib/transforms.py:22-38    torchvision Resize / ToTensor / Normalize
//   wsinsight/modellib/run_inference.py:288-299,543  default collate + H2D copy
//
// The source window of each patch is staged into shared memory by TMA
// (cp.async.bulk.tensor, zero fill outside the slide == the reference's padding); the
// resample follows Pillow's Resample.c in 22-bit fixed point with a uint8 intermediate
// after the horizontal pass, so the u8 result is bit-exact with PIL; x/255, (x-mean)/std
// are done in fp32 exactly like torchvision and rounded once to bf16 (RNE).
#include "common.cuh"
#include "../../include/wsi_b200.h"
#include <math.h>
#include <vector>

namespace {

constexpr int PRECISION_BITS = 32 - 8 - 2;
constexpr int PP_THREADS = 256;
constexpr int PP_BOX_W = 256;        // source pixels per TMA box row (1024 B)
constexpr int PP_MAX_INTER_ROWS = 64;

struct Taps {
  std::vector<int> lo, cnt, coef;
  int ksize = 0;
};

// Pillow precompute_coeffs() for the BILINEAR (triangle, support 1) filter + normalize_coeffs_8bpc()
Taps make_taps(int in_size, int out_size) {
  Taps t;
  double scale = (double)in_size / (double)out_size;
  double filterscale = scale < 1.0 ? 1.0 : scale;
  double support = 1.0 * filterscale;
  int ksize = (int)ceil(support) * 2 + 1;
  t.ksize = ksize;
  t.lo.resize(out_size);
  t.cnt.resize(out_size);
  t.coef.assign((size_t)out_size * ksize, 0);
  std::vector<double> w(ksize);
  double ss = 1.0 / filterscale;
  for (int xx = 0; xx < out_size; ++xx) {
    double center = (xx + 0.5) * scale;
    int lo = (int)(center - support + 0.5);
    if (lo < 0) lo = 0;
    int hi = (int)(center + support + 0.5);
    if (hi > in_size) hi = in_size;
    int n = hi - lo;
    double ww = 0.0;
    for (int x = 0; x < n; ++x) {
      double v = (x + lo - center + 0.5) * ss;
      if (v < 0) v = -v;
      double wv = v < 1.0 ? 1.0 - v : 0.0;
      w[x] = wv;
      ww += wv;
    }
    for (int x = 0; x < n; ++x) {
      if (ww != 0.0) w[x] /= ww;
      double q = w[x] * (double)(1 << PRECISION_BITS);
      t.coef[(size_t)xx * ksize + x] = (int)(w[x] < 0 ? -0.5 + q : 0.5 + q);
    }
    t.lo[xx] = lo;
    t.cnt[xx] = n;
  }
  return t;
}

struct Plan {
  int ps, S, ksize;
  int TR, nbands, ngroups, G;
  int box_w, box_h;               // TMA box (pixels, rows)
  int max_src_rows;               // intermediate rows per band
  int out_H, out_W, pad_top, pad_left;
  float mean[3], stdv[3];
  int has_norm;
  long long plane_stride_px = 0;  // > 0: row-parity plane output layout (conv_stem.cuh)
  int capacity = 0;               // > 0: patches the dense output buffer holds (wsi_preproc_plan_set_capacity)
  int channels = 4;               // 4 | 8 bf16 per output pixel (channels >= 3 are zero)
  int *d_lo, *d_cnt, *d_coef;     // taps (same table for both axes: square patches)
  uint16_t* d_lut = nullptr;      // [3][256] bf16 bits of ToTensor + Normalize applied to a u8 sample
  size_t smem_bytes;
  CUtensorMap tmap;
};

struct PPArgs {
  const int* lo;
  const int* cnt;
  const int* coef;
  const uint16_t* lut;
  const int32_t* coords;   // (n,2) top-left x,y
  __nv_bfloat16* out;
  uint32_t* out_u8;        // debug/parity: resized RGB0 u8 [n][S][S] (out is ignored when set)
  int ps, S, ksize, TR, G, ngroups, box_w, box_h, max_src_rows;
  int out_H, out_W, pad_top, pad_left;
  long long plane_stride_px;   // > 0: write two row-parity planes [2][maxB][P][2P] of 4-channel pixels
  int channels;                // 4: 8-byte pixels, 8: 16-byte pixels
};

// Pillow clips (sum >> PRECISION_BITS) to [0, 255].  For the triangle filter every coefficient is >= 0 and a row sums
// to 2^22 + at most ksize/2 rounding units, so HALF + sum(c * p) lies in [2^21, 255 * 2^22 + 255 * ksize/2 + 2^21]
// < 256 * 2^22 for any ksize < 16384 (checked when the plan is built): the clip can never fire and is left out
// (six min/max instructions per pixel in an issue-bound kernel).
__device__ __forceinline__ uint32_t clip8(int v) { return (uint32_t)v >> PRECISION_BITS; }
// bytes 0..2 of an RGBx word, zero extended (one PRMT each instead of shift + mask)
__device__ __forceinline__ int byte0(uint32_t p) { return (int)(p & 0xffu); }
__device__ __forceinline__ int byte1(uint32_t p) { return (int)__byte_perm(p, 0u, 0x4441u); }
__device__ __forceinline__ int byte2(uint32_t p) { return (int)__byte_perm(p, 0u, 0x4442u); }

// grid: (n_patches, nbands); one CTA = TR output rows x S output columns of one patch.
// One thread owns one output column in both passes: its horizontal taps sit in registers (KS > 0) and the
// vertical taps of a row are warp-uniform shared-memory broadcasts, so the inner loops are LDS + IMAD only.
// ToTensor + Normalize + the bf16 rounding are a function of the u8 sample alone: a 3 x 256 table built on the
// host with the same IEEE fp32 operations torchvision performs (x / 255, (x - mean) / std, one RNE).
template <int KS>
__global__ void __launch_bounds__(PP_THREADS)
    k_preprocess(const __grid_constant__ CUtensorMap tmap, PPArgs a) {
  extern __shared__ __align__(128) uint8_t smem[];
  uint32_t* s_src = reinterpret_cast<uint32_t*>(smem);                          // box_h x box_w RGBA
  uint32_t* s_int = s_src + (size_t)a.box_h * a.box_w;                          // max_src_rows x S (RGBx u8)
  int* s_lo = reinterpret_cast<int*>(s_int + (size_t)a.max_src_rows * a.S);     // S
  int* s_cnt = s_lo + a.S;                                                      // S
  int* s_coef = s_cnt + a.S;                                                    // S x ksize
  uint16_t* s_lut = reinterpret_cast<uint16_t*>(s_coef + (size_t)a.S * a.ksize);   // 3 x 256 bf16 bit patterns
  __shared__ __align__(8) uint64_t s_bar;

  const int patch = blockIdx.x;
  const int r0 = blockIdx.y * a.TR;
  const int r1 = min(r0 + a.TR, a.S);
  const int px = a.coords[2 * patch + 0], py = a.coords[2 * patch + 1];

  if (threadIdx.x == 0) {
    ptx::mbar_init(&s_bar, 1);
    ptx::fence_barrier_init();
  }
  for (int i = threadIdx.x; i < a.S; i += PP_THREADS) {
    s_lo[i] = a.lo[i];
    s_cnt[i] = a.cnt[i];
  }
  for (int i = threadIdx.x; i < a.S * a.ksize; i += PP_THREADS) s_coef[i] = a.coef[i];
  for (int i = threadIdx.x; i < 768; i += PP_THREADS) s_lut[i] = a.lut[i];
  __syncthreads();
  const int y_lo = s_lo[r0];
  const int y_hi = s_lo[r1 - 1] + s_cnt[r1 - 1];
  const int nrows = y_hi - y_lo;
  uint32_t phase = 0;
  constexpr int HALF = 1 << (PRECISION_BITS - 1);

  // horizontal pass: source rows [y_lo, y_hi) x output columns, chunked by TMA boxes
  for (int yc = 0; yc < nrows; yc += a.box_h) {
    const int rows_here = min(a.box_h, nrows - yc);
    for (int g = 0; g < a.ngroups; ++g) {
      const int xx0 = g * a.G, xx1 = min(xx0 + a.G, a.S);
      const int xs0 = s_lo[xx0];
      // TMA needs the box start 16-byte aligned in global memory: align x down to 4 pixels
      const int bx = (px + xs0) & ~3;
      const int shift = (px + xs0) - bx;
      if (threadIdx.x == 0) {
        ptx::mbar_arrive_expect_tx(&s_bar, (uint32_t)(a.box_h * a.box_w * 4));
        ptx::tma_load_2d(s_src, &tmap, &s_bar, bx, py + y_lo + yc);
      }
      ptx::mbar_wait(&s_bar, phase);
      phase ^= 1;
      for (int xx = xx0 + threadIdx.x; xx < xx1; xx += PP_THREADS) {
        const uint32_t* src = s_src + (s_lo[xx] - xs0 + shift);
        uint32_t* dst = s_int + (size_t)yc * a.S + xx;
        if (KS > 0) {
          // taps beyond cnt carry zero coefficients (the table is zero padded); the pixels they touch are
          // inside the shared-memory allocation, whatever they hold
          int c[KS > 0 ? KS : 1];
#pragma unroll
          for (int t = 0; t < KS; ++t) c[t] = s_coef[xx * KS + t];
#pragma unroll 4
          for (int rr = 0; rr < rows_here; ++rr) {
            int s0 = HALF, s1 = HALF, s2 = HALF;
#pragma unroll
            for (int t = 0; t < KS; ++t) {
              const uint32_t p = src[rr * a.box_w + t];
              s0 += byte0(p) * c[t];
              s1 += byte1(p) * c[t];
              s2 += byte2(p) * c[t];
            }
            dst[rr * a.S] = clip8(s0) | (clip8(s1) << 8) | (clip8(s2) << 16);
          }
        } else {
          const int n = s_cnt[xx];
          const int* k = s_coef + (size_t)xx * a.ksize;
          for (int rr = 0; rr < rows_here; ++rr) {
            int s0 = HALF, s1 = HALF, s2 = HALF;
            for (int t = 0; t < n; ++t) {
              const uint32_t p = src[rr * a.box_w + t];
              const int c = k[t];
              s0 += byte0(p) * c;
              s1 += byte1(p) * c;
              s2 += byte2(p) * c;
            }
            dst[rr * a.S] = clip8(s0) | (clip8(s1) << 8) | (clip8(s2) << 16);
          }
        }
      }
      __syncthreads();   // s_src is reused by the next box; s_int rows complete
    }
  }

  // vertical pass + ToTensor + Normalize + bf16 NHWC4 store
  const int P = a.out_H >> 1;
  for (int xx = threadIdx.x; xx < a.S; xx += PP_THREADS) {
    // destination index of (row r0, column xx), then advanced row by row: in the row-parity-plane layout an even
    // stored row y is followed by y + 1 one plane further, an odd one by the next row of plane 0
    long long o;
    {
      const int y = a.pad_top + r0;
      if (a.plane_stride_px > 0)
        o = (long long)(y & 1) * a.plane_stride_px + ((long long)patch * P + (y >> 1)) * a.out_W + (a.pad_left + xx);
      else
        o = ((long long)patch * a.out_H + y) * a.out_W + (a.pad_left + xx);
    }
    const long long step_even = a.plane_stride_px > 0 ? (long long)a.plane_stride_px : (long long)a.out_W;
    const long long step_odd = a.plane_stride_px > 0 ? (long long)a.out_W - a.plane_stride_px : (long long)a.out_W;
    int ypar = (a.pad_top + r0) & 1;
    for (int yy = r0; yy < r1; ++yy, o += ypar ? step_odd : step_even, ypar ^= 1) {
      const int n = s_cnt[yy];                                   // warp-uniform
      const uint32_t* col = s_int + (size_t)(s_lo[yy] - y_lo) * a.S + xx;
      const int* k = s_coef + (size_t)yy * a.ksize;
      int s0 = HALF, s1 = HALF, s2 = HALF;
      if (KS > 0) {
        // as in the horizontal pass: taps beyond cnt carry zero coefficients and the rows they touch lie inside the
        // shared-memory allocation (at most KS - 1 rows past the staged ones, into the tables that follow s_int)
#pragma unroll
        for (int t = 0; t < KS; ++t) {
          const uint32_t p = col[t * a.S];
          const int c = k[t];
          s0 += byte0(p) * c;
          s1 += byte1(p) * c;
          s2 += byte2(p) * c;
        }
      } else {
        for (int t = 0; t < n; ++t) {
          const uint32_t p = col[t * a.S];
          const int c = k[t];
          s0 += byte0(p) * c;
          s1 += byte1(p) * c;
          s2 += byte2(p) * c;
        }
      }
      const uint32_t u0 = clip8(s0), u1 = clip8(s1), u2 = clip8(s2);
      if (a.out_u8) {
        a.out_u8[((size_t)patch * a.S + yy) * a.S + xx] = u0 | (u1 << 8) | (u2 << 16);
        continue;
      }
      uint2 v;
      v.x = (uint32_t)s_lut[u0] | ((uint32_t)s_lut[256 + u1] << 16);
      v.y = (uint32_t)s_lut[512 + u2];
      if (a.channels == 8)
        reinterpret_cast<uint4*>(a.out)[o] = make_uint4(v.x, v.y, 0u, 0u);
      else
        reinterpret_cast<uint2*>(a.out)[o] = v;
    }
  }
}

}  // namespace

extern "C" {

int wsi_preproc_plan_create(const uint8_t* d_slide_rgba, long long slide_w, long long slide_h,
                            long long pitch_bytes, int patch_size, int out_size, const float* mean,
                            const float* stdv, int out_H, int out_W, int pad_top, int pad_left,
                            void** plan_out) {
  WSI_ARG_CHECK(d_slide_rgba && plan_out, "wsi_preproc_plan_create: null pointer");
  WSI_ARG_CHECK(((uintptr_t)d_slide_rgba & 15) == 0 && (pitch_bytes & 15) == 0 && pitch_bytes >= slide_w * 4,
                "wsi_preproc_plan_create: slide base/pitch must be 16-byte aligned RGBA");
  WSI_ARG_CHECK(patch_size >= 1 && out_size >= 1 && out_size <= 1024, "wsi_preproc_plan_create: bad sizes");
  WSI_ARG_CHECK(out_H >= pad_top + out_size && out_W >= pad_left + out_size && pad_top >= 0 && pad_left >= 0,
                "wsi_preproc_plan_create: output geometry too small");
  Plan* p = new Plan();
  p->ps = patch_size;
  p->S = out_size;
  p->out_H = out_H;
  p->out_W = out_W;
  p->pad_top = pad_top;
  p->pad_left = pad_left;
  p->has_norm = (mean && stdv) ? 1 : 0;
  for (int c = 0; c < 3; ++c) {
    p->mean[c] = mean ? mean[c] : 0.f;
    p->stdv[c] = stdv ? stdv[c] : 1.f;
  }
  Taps t = make_taps(patch_size, out_size);
  p->ksize = t.ksize;
  // the kernels rely on the clip to [0, 255] being a no-op (see clip8): non-negative taps, row sums ~ 2^22
  for (int i = 0; i < out_size; ++i) {
    long long sum = 0;
    for (int k = 0; k < t.ksize; ++k) {
      const int c = t.coef[(size_t)i * t.ksize + k];
      if (c < 0) {
        delete p;
        wsi_set_error("wsi_preproc_plan_create: negative resampling coefficient");
        return WSI_ERR_UNSUPPORTED;
      }
      sum += c;
    }
    if (sum > (1LL << PRECISION_BITS) + 8192) {
      delete p;
      wsi_set_error("wsi_preproc_plan_create: resampling coefficients of output %d sum to %lld", i, sum);
      return WSI_ERR_UNSUPPORTED;
    }
  }
  const int S = out_size;
  // rows per band so that the u8 intermediate stays <= PP_MAX_INTER_ROWS source rows
  int TR = 16;
  for (;;) {
    int worst = 0;
    for (int r0 = 0; r0 < S; r0 += TR) {
      int r1 = r0 + TR < S ? r0 + TR : S;
      int rows = t.lo[r1 - 1] + t.cnt[r1 - 1] - t.lo[r0];
      if (rows > worst) worst = rows;
    }
    p->max_src_rows = worst;
    if (worst <= PP_MAX_INTER_ROWS || TR == 1) break;
    TR = TR / 2;
  }
  p->TR = TR;
  p->nbands = (S + TR - 1) / TR;
  // output-column groups whose source span fits one TMA box row
  int box_w = patch_size + 8 + 3;   // +3: the box start is aligned down to 4 pixels (16 B) for TMA
  box_w = (box_w + 3) / 4 * 4;
  if (box_w > PP_BOX_W) box_w = PP_BOX_W;
  int G = S;
  for (;;) {
    int worst = 0;
    for (int x0 = 0; x0 < S; x0 += G) {
      int x1 = x0 + G < S ? x0 + G : S;
      int span = t.lo[x1 - 1] + t.cnt[x1 - 1] - t.lo[x0];
      if (span > worst) worst = span;
    }
    if (worst + 3 <= box_w || G == 1) {
      if (worst + 3 > box_w) {
        delete p;
        wsi_set_error("wsi_preproc_plan_create: filter support %d exceeds TMA box %d", worst, box_w);
        return WSI_ERR_UNSUPPORTED;
      }
      break;
    }
    G = (G + 1) / 2;
  }
  p->G = G;
  p->ngroups = (S + G - 1) / G;
  p->box_w = box_w;
  int box_h = (32 * 1024) / (box_w * 4);
  if (box_h > p->max_src_rows) box_h = p->max_src_rows;
  if (box_h > 256) box_h = 256;
  if (box_h < 1) box_h = 1;
  p->box_h = box_h;
  p->smem_bytes = (size_t)box_h * box_w * 4 + (size_t)p->max_src_rows * S * 4 + (size_t)S * 4 * (2 + t.ksize) + 768 * 2;
  if (p->smem_bytes > 200 * 1024) {
    delete p;
    wsi_set_error("wsi_preproc_plan_create: shared memory budget exceeded (%zu)", p->smem_bytes);
    return WSI_ERR_UNSUPPORTED;
  }
  size_t nb = (size_t)S * sizeof(int);
  if (cudaMalloc(&p->d_lo, nb) != cudaSuccess || cudaMalloc(&p->d_cnt, nb) != cudaSuccess ||
      cudaMalloc(&p->d_coef, nb * t.ksize) != cudaSuccess || cudaMalloc(&p->d_lut, 768 * sizeof(uint16_t)) != cudaSuccess) {
    delete p;
    wsi_set_error("wsi_preproc_plan_create: cudaMalloc failed");
    return WSI_ERR_NOMEM;
  }
  cudaMemcpy(p->d_lo, t.lo.data(), nb, cudaMemcpyHostToDevice);
  cudaMemcpy(p->d_cnt, t.cnt.data(), nb, cudaMemcpyHostToDevice);
  cudaMemcpy(p->d_coef, t.coef.data(), nb * t.ksize, cudaMemcpyHostToDevice);
  {
    // torchvision: ToTensor = u8 / 255 (fp32), Normalize = (x - mean) / std (fp32); then one RNE to bf16
    uint16_t lut[768];
    for (int c = 0; c < 3; ++c)
      for (int v = 0; v < 256; ++v) {
        volatile float f = (float)v / 255.0f;
        if (p->has_norm) {
          volatile float d = f - p->mean[c];
          f = d / p->stdv[c];
        }
        __nv_bfloat16 b = __float2bfloat16_rn(f);
        lut[c * 256 + v] = *reinterpret_cast<uint16_t*>(&b);
      }
    cudaMemcpy(p->d_lut, lut, sizeof(lut), cudaMemcpyHostToDevice);
  }
  uint64_t dims[2] = {(uint64_t)slide_w, (uint64_t)slide_h};
  uint64_t strides[1] = {(uint64_t)pitch_bytes};
  uint32_t box[2] = {(uint32_t)box_w, (uint32_t)box_h};
  uint32_t es[2] = {1, 1};
  int rc = wsi_make_tmap(&p->tmap, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, d_slide_rgba, dims, strides, box, es,
                         CU_TENSOR_MAP_SWIZZLE_NONE);
  if (rc) {
    cudaFree(p->d_lo);
    cudaFree(p->d_cnt);
    cudaFree(p->d_coef);
    cudaFree(p->d_lut);
    delete p;
    return rc;
  }
  cudaError_t e = cudaFuncSetAttribute(k_preprocess<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k_preprocess<5>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k_preprocess<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  if (e != cudaSuccess) {
    wsi_set_error("wsi_preproc_plan_create: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    return WSI_ERR_CUDA;
  }
  *plan_out = p;
  return WSI_OK;
}

static int preproc_launch(void* plan, const int32_t* d_coords_xy, int n, void* d_out_bf16, uint8_t* d_out_u8,
                          void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  WSI_ARG_CHECK(plan && d_coords_xy && (d_out_bf16 || d_out_u8), "wsi_preproc_run: null pointer");
  if (n <= 0) return WSI_OK;
  Plan* p = (Plan*)plan;
  // parity-plane layout: plane 0 holds at most plane_stride_px / ((out_H / 2) * out_W) patches before it runs into plane 1
  WSI_ARG_CHECK(d_out_u8 || p->plane_stride_px == 0 ||
                    (long long)n * (p->out_H / 2) * p->out_W <= p->plane_stride_px,
                "wsi_preproc_run: %d patches exceed the capacity of the parity-plane buffer (%lld)", n,
                p->plane_stride_px / ((long long)(p->out_H / 2) * p->out_W));
  WSI_ARG_CHECK(d_out_u8 || p->capacity <= 0 || n <= p->capacity,
                "wsi_preproc_run: %d patches exceed the output capacity %d", n, p->capacity);
  PPArgs a;
  a.out_u8 = (uint32_t*)d_out_u8;
  a.lo = p->d_lo;
  a.cnt = p->d_cnt;
  a.coef = p->d_coef;
  a.coords = d_coords_xy;
  a.out = (__nv_bfloat16*)d_out_bf16;
  a.ps = p->ps;
  a.S = p->S;
  a.ksize = p->ksize;
  a.TR = p->TR;
  a.G = p->G;
  a.ngroups = p->ngroups;
  a.box_w = p->box_w;
  a.box_h = p->box_h;
  a.max_src_rows = p->max_src_rows;
  a.out_H = p->out_H;
  a.out_W = p->out_W;
  a.pad_top = p->pad_top;
  a.pad_left = p->pad_left;
  a.lut = p->d_lut;
  a.plane_stride_px = p->plane_stride_px;
  a.channels = p->channels;
  dim3 grid(n, p->nbands);
  if (p->ksize == 3)
    k_preprocess<3><<<grid, PP_THREADS, p->smem_bytes, stream>>>(p->tmap, a);
  else if (p->ksize == 5)
    k_preprocess<5><<<grid, PP_THREADS, p->smem_bytes, stream>>>(p->tmap, a);
  else
    k_preprocess<0><<<grid, PP_THREADS, p->smem_bytes, stream>>>(p->tmap, a);
  wsi_count_launch(1);
  WSI_CUDA_CHECK(cudaGetLastError());
  return WSI_OK;
}

int wsi_preproc_run(void* plan, const int32_t* d_coords_xy, int n, void* d_out_bf16, void* stream_) {
  WSI_ARG_CHECK(d_out_bf16, "wsi_preproc_run: null output");
  return preproc_launch(plan, d_coords_xy, n, d_out_bf16, nullptr, stream_);
}

int wsi_preproc_run_u8(void* plan, const int32_t* d_coords_xy, int n, uint8_t* d_out_u8, void* stream_) {
  WSI_ARG_CHECK(d_out_u8, "wsi_preproc_run_u8: null output");
  return preproc_launch(plan, d_coords_xy, n, nullptr, d_out_u8, stream_);
}

int wsi_preproc_plan_set_parity_layout(void* plan, long long plane_stride_px) {
  WSI_ARG_CHECK(plan && plane_stride_px >= 0, "wsi_preproc_plan_set_parity_layout: bad arguments");
  Plan* p = (Plan*)plan;
  WSI_ARG_CHECK(plane_stride_px == 0 || (p->out_H == p->out_W && (p->out_H & 1) == 0),
                "wsi_preproc_plan_set_parity_layout: the padded output must be square with an even side");
  p->plane_stride_px = plane_stride_px;
  return WSI_OK;
}

int wsi_preproc_plan_set_capacity(void* plan, int max_patches) {
  WSI_ARG_CHECK(plan && max_patches >= 0, "wsi_preproc_plan_set_capacity: bad arguments");
  ((Plan*)plan)->capacity = max_patches;
  return WSI_OK;
}

int wsi_preproc_plan_set_channels(void* plan, int channels) {
  WSI_ARG_CHECK(plan && (channels == 4 || channels == 8), "wsi_preproc_plan_set_channels: channels must be 4 or 8");
  Plan* p = (Plan*)plan;
  WSI_ARG_CHECK(channels == 4 || p->plane_stride_px == 0, "wsi_preproc_plan_set_channels: the parity layout has 4 channels");
  p->channels = channels;
  return WSI_OK;
}

int wsi_preproc_plan_info(void* plan, int* info8) {
  WSI_ARG_CHECK(plan && info8, "wsi_preproc_plan_info: null pointer");
  Plan* p = (Plan*)plan;
  info8[0] = p->TR; info8[1] = p->nbands; info8[2] = p->G; info8[3] = p->ngroups;
  info8[4] = p->box_w; info8[5] = p->box_h; info8[6] = p->max_src_rows; info8[7] = (int)p->smem_bytes;
  return WSI_OK;
}

void wsi_preproc_plan_destroy(void* plan) {
  if (!plan) return;
  Plan* p = (Plan*)plan;
  cudaFree(p->d_lo);
  cudaFree(p->d_cnt);
  cudaFree(p->d_coef);
  cudaFree(p->d_lut);
  delete p;
}

}  // extern "C"
