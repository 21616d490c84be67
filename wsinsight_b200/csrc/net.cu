// K4: network runtime -- a small layer-graph executor over the tcgen05 conv kernel.
//
// Replaces the TorchScript forward + CPU softmax of the reference classification loop
// (wsinsight/modellib/run_inference.py:537-560) for the zoo CNNs (ResNet34, InceptionV4):
// the Python side walks the (frozen) TorchScript graph, hands folded conv weights/biases
// to wsi_net_add_*, and every forward runs only kernels from this library.
#include "conv_flat.cuh"
#include "conv_flat3.cuh"
#include "conv_stem.cuh"
#include "../../include/wsi_b200.h"
#include <math.h>
#include <stdlib.h>
#include <vector>

namespace {

struct Tensor {
  int H, W, C;
  __nv_bfloat16* ptr = nullptr;   // [max_batch][H+2*pad][W+2*pad][C]; nullptr for the external input
  bool external = false;
  int pad = 0;                    // one-pixel zero border (tensors touched by the flat 3x3 conv kernel)
  int xh = 0, xw = 0;             // extra junk rows / columns after the image (output of the flat stem kernel)
  bool lock_geom = false;         // border / junk geometry fixed by the producing kernel (parity stem output)
  bool elided = false;            // never materialised: the stem output when the max pooling is fused into the stem
  int Hp() const { return H + 2 * pad + xh; }
  int Wp() const { return W + 2 * pad + xw; }
  size_t interior_offset() const { return ((size_t)pad * Wp() + pad) * C; }   // elements to pixel (0,0)
};

enum OpKind { OP_CONV, OP_MAXPOOL, OP_AVGPOOL, OP_HEAD };

struct Op {
  OpKind kind;
  int in, out, res;
  // conv
  convk::Args ca;
  convf::Args fa;
  convs::Args sa;
  bool flat = false;              // 3x3/s1/p1 over padded tensors -> convf::k_conv_flat
  bool flat3 = false;             // ... with Cout = 64 and short rows -> convf3::k_conv_flat3 (kw-split accumulators)
  convf3::Args f3;
  int fused_pool = -1;            // stem: index of the 3x3/2 max-pool op fused into its epilogue
  bool skip = false;              // pool op executed by the stem kernel
  int cin_real = 0;               // input channels that carry a non-zero weight (<= Cin)
  int max_pairs = 0;              // co-resident 2-CTA clusters of the per-tap kernel (queried at first launch)
  int pt_ct = 0, pt_wt = 0, pt_cols = 0, pt_R = 0, pt_per = 1;   // 3x3 pool through the TMA ring: tile (channel vectors, output columns), box columns, depth
  CUtensorMap tmA, tmB, tmO, tmR, tmO32;
  __nv_bfloat16* d_w = nullptr;
  float* d_bias = nullptr;
  std::vector<__nv_bfloat16> h_w;
  std::vector<float> h_bias;
  int Cin, Cout, KH, KW, stride, pad_h, pad_w, relu, out_c_off, is_stem;
  int in_c_off = 0;               // first input channel read (slice of a wider tensor)
  int relu_cols = -1;             // conv: ReLU only on output channels < relu_cols (-1: all, when relu is set)
  int res_c_off = 0;              // residual = channels [res_c_off, res_c_off + Cout) of tensor `res`
  int center_cols = -1;           // output channels >= center_cols carry weights at the centre tap only (-1: none)
  int pC = 0;                     // pool: channels pooled (slice [in_c_off, in_c_off + pC) of the input tensor)
  int Ktot;
  // pool
  int pk, pstride, ppad, count_include_pad;
  int prelu = 0;                  // average pool followed by + bias, ReLU (a 1x1 conv commuted in front of the pool)
  float* d_pbias = nullptr;
  std::vector<float> h_pbias;
  // head
  int K;
  float* d_fcw = nullptr;
  float* d_fcb = nullptr;
  std::vector<float> h_fcw, h_fcb;
  double flops_per_image = 0;
};

struct Net {
  int max_batch;
  bool finalized = false;
  std::vector<Tensor> tensors;
  std::vector<Op> ops;
  int num_sms = 148;
  bool flat_enabled = true;       // WSI_B200_NO_FLAT=1 forces every conv onto the per-tap TMA kernel
  bool cluster_enabled = true;    // WSI_B200_NO_CLUSTER=1: no CTA pairs / weight multicast in the per-tap kernel
  bool no_flat_pair = false;      // WSI_B200_NO_FLAT_PAIR=1: the flat kernel stays single-CTA
  bool res_smem_enabled = true;   // WSI_B200_NO_RES_SMEM=1: flat-kernel residual through registers (scattered loads)
  bool pool_ring = true;          // WSI_B200_NO_POOL_RING=1: 3x3 pools by the register-ring kernel (direct global loads)
  bool legacy_epi = false;        // WSI_B200_LEGACY_EPI=1: per-thread global stores for N tiles that are not multiples of 64
  bool stem_pool_enabled = true;  // WSI_B200_NO_STEM_POOL=1: stem conv and max pooling stay two kernels
  bool pdl_enabled = true;        // WSI_B200_NO_PDL=1: conv kernels launched without programmatic dependent launch
  bool kw3_enabled = false;       // WSI_B200_KW3=1: Cout = 64 flat convs on the kw-split kernel (measured slower, see conv_flat3.cuh)
  int launches_per_forward = 0;
  // optional per-op CUDA-event timing (bench roofline): ring of `prof_slots` event sets
  int prof_slots = 0;
  long long prof_forwards = 0;
  std::vector<cudaEvent_t> prof_ev;   // [prof_slots][ops + 1]
};

// ---------------------------------------------------------------------------
// pooling / head kernels (NHWC bf16, 8 channels = 16 B per thread)
// ---------------------------------------------------------------------------
__device__ __forceinline__ float bflo(uint32_t v) { return __uint_as_float(v << 16); }
__device__ __forceinline__ float bfhi(uint32_t v) { return __uint_as_float(v & 0xffff0000u); }

// (iHp, iWp, ip) / (oHp, oWp, op): stored rows, stored columns and border of the input / output tensor; `in`/`out` point at the buffer start.
// One CTA per output row (n, ho); threadIdx.x walks the 16-byte channel vectors, threadIdx.y the output columns:
// no integer division on the data path, 32-bit offsets inside an image row, the k x k window re-read from L1.
template <bool IS_MAX>
__global__ void __launch_bounds__(256)
    k_pool(const __nv_bfloat16* __restrict__ in, int H, int W, int C, int k, int s, int pad, int count_include_pad,
           int Ho, int Wo, __nv_bfloat16* __restrict__ out, int out_C, int out_c_off, int iHp, int iWp, int ip, int oHp,
           int oWp, int op) {
  const int cv = C >> 3;
  const int n = blockIdx.x / Ho, ho = blockIdx.x - n * Ho;
  const int y0 = max(ho * s - pad, 0), y1 = min(ho * s - pad + k, H);
  const uint4* img = reinterpret_cast<const uint4*>(in + (size_t)n * iHp * iWp * C);
  uint4* orow = reinterpret_cast<uint4*>(out + (((size_t)n * oHp + ho + op) * oWp + op) * out_C + out_c_off);
  const int ocv = out_C >> 3;
  for (int wo = threadIdx.y; wo < Wo; wo += blockDim.y) {
    const int x0 = max(wo * s - pad, 0), x1 = min(wo * s - pad + k, W);
    const float d = count_include_pad ? (float)(k * k) : (float)((y1 - y0) * (x1 - x0));
    for (int c8 = threadIdx.x; c8 < cv; c8 += blockDim.x) {
      float m[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) m[i] = IS_MAX ? -INFINITY : 0.f;
      for (int y = y0; y < y1; ++y) {
        const uint4* row = img + ((size_t)(y + ip) * iWp + ip) * cv + c8;
        for (int x = x0; x < x1; ++x) {
          const uint4 v = __ldg(row + (size_t)x * cv);
          if (IS_MAX) {
            m[0] = fmaxf(m[0], bflo(v.x)); m[1] = fmaxf(m[1], bfhi(v.x));
            m[2] = fmaxf(m[2], bflo(v.y)); m[3] = fmaxf(m[3], bfhi(v.y));
            m[4] = fmaxf(m[4], bflo(v.z)); m[5] = fmaxf(m[5], bfhi(v.z));
            m[6] = fmaxf(m[6], bflo(v.w)); m[7] = fmaxf(m[7], bfhi(v.w));
          } else {
            m[0] += bflo(v.x); m[1] += bfhi(v.x); m[2] += bflo(v.y); m[3] += bfhi(v.y);
            m[4] += bflo(v.z); m[5] += bfhi(v.z); m[6] += bflo(v.w); m[7] += bfhi(v.w);
          }
        }
      }
      if (!IS_MAX) {
#pragma unroll
        for (int i = 0; i < 8; ++i) m[i] = m[i] / d;
      }
      uint4 o;
      o.x = convk::pack_bf16(m[0], m[1]);
      o.y = convk::pack_bf16(m[2], m[3]);
      o.z = convk::pack_bf16(m[4], m[5]);
      o.w = convk::pack_bf16(m[6], m[7]);
      orow[(size_t)wo * ocv + c8] = o;
    }
  }
}

// Separable sliding-window pooling for the 3x3 windows of the zoo CNNs (stride 1 or 2).  The generic kernel above
// re-reads every input vector k*k times through L1/L2 (ncu/bench: 1 TB/s of DRAM traffic, i.e. ~9 TB/s of L2->SM
// traffic on the Inception-A average pools).  Here a thread owns one (channel vector, output column) and walks
// DOWN the image: each input row is reduced horizontally once (K loads, neighbours shared through L1) into a
// K-deep register ring, the vertical reduction is register-only, and consecutive output rows reuse K - S rows.
__device__ __forceinline__ uint32_t hmax2_bf16(uint32_t a, uint32_t b) {
  __nv_bfloat162 r = __hmax2(*reinterpret_cast<__nv_bfloat162*>(&a), *reinterpret_cast<__nv_bfloat162*>(&b));
  return *reinterpret_cast<uint32_t*>(&r);
}

// Row accumulator of the sliding-window pool: packed bf16x2 for max (exact, 4 registers per row), fp32 for sums.
template <bool IS_MAX>
struct PoolRow;
template <>
struct PoolRow<true> {
  uint4 v;
  __device__ __forceinline__ void clear() { v = make_uint4(0xff80ff80u, 0xff80ff80u, 0xff80ff80u, 0xff80ff80u); }   // -inf
  __device__ __forceinline__ void add(const uint4& x) {
    v.x = hmax2_bf16(v.x, x.x); v.y = hmax2_bf16(v.y, x.y); v.z = hmax2_bf16(v.z, x.z); v.w = hmax2_bf16(v.w, x.w);
  }
  __device__ __forceinline__ void merge(const PoolRow& o) { add(o.v); }
  __device__ __forceinline__ void set(const uint4& x) { v = x; }
  static __device__ __forceinline__ uint4 neutral() { return make_uint4(0xff80ff80u, 0xff80ff80u, 0xff80ff80u, 0xff80ff80u); }
  __device__ __forceinline__ uint4 result(float) const { return v; }
  __device__ __forceinline__ uint4 result_bias(float, const float (&)[8], int) const { return v; }   // max: no bias
};
template <>
struct PoolRow<false> {
  float f[8];
  __device__ __forceinline__ void clear() {
#pragma unroll
    for (int i = 0; i < 8; ++i) f[i] = 0.f;
  }
  __device__ __forceinline__ void add(const uint4& x) {
    f[0] += bflo(x.x); f[1] += bfhi(x.x); f[2] += bflo(x.y); f[3] += bfhi(x.y);
    f[4] += bflo(x.z); f[5] += bfhi(x.z); f[6] += bflo(x.w); f[7] += bfhi(x.w);
  }
  __device__ __forceinline__ void merge(const PoolRow& o) {
#pragma unroll
    for (int i = 0; i < 8; ++i) f[i] += o.f[i];
  }
  __device__ __forceinline__ void set(const uint4& x) {
    f[0] = bflo(x.x); f[1] = bfhi(x.x); f[2] = bflo(x.y); f[3] = bfhi(x.y);
    f[4] = bflo(x.z); f[5] = bfhi(x.z); f[6] = bflo(x.w); f[7] = bfhi(x.w);
  }
  static __device__ __forceinline__ uint4 neutral() { return make_uint4(0u, 0u, 0u, 0u); }
  // one reciprocal per output instead of eight IEEE divisions (ncu: the divisions made the kernel issue-bound,
  // 76 % issue-slot utilisation at 1.6 TB/s); <= 1 ulp of fp32 before the bf16 rounding
  __device__ __forceinline__ uint4 result(float rcp) const {
    return make_uint4(convk::pack_bf16(f[0] * rcp, f[1] * rcp), convk::pack_bf16(f[2] * rcp, f[3] * rcp),
                      convk::pack_bf16(f[4] * rcp, f[5] * rcp), convk::pack_bf16(f[6] * rcp, f[7] * rcp));
  }
  // average, then + bias and optional ReLU (the epilogue of a 1x1 conv that was moved in front of the pool)
  __device__ __forceinline__ uint4 result_bias(float rcp, const float (&b)[8], int relu) const {
    float v[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      v[i] = fmaf(f[i], rcp, b[i]);
      if (relu) v[i] = fmaxf(v[i], 0.f);
    }
    return make_uint4(convk::pack_bf16(v[0], v[1]), convk::pack_bf16(v[2], v[3]), convk::pack_bf16(v[4], v[5]),
                      convk::pack_bf16(v[6], v[7]));
  }
};

template <bool IS_MAX, int K, int S>
__global__ void __launch_bounds__(256, 6)
    k_pool_sep(const __nv_bfloat16* __restrict__ in, int H, int W, int C, int pad, int count_include_pad, int Ho, int Wo,
               __nv_bfloat16* __restrict__ out, int out_C, int out_c_off, int iHp, int iWp, int ip, int oHp, int oWp,
               int op, int band, int nbands, const float* __restrict__ pbias, int prelu) {
  // threads walk the (output column, 16-byte channel vector) plane of a row in flat order: no idle lanes whatever
  // Wo and C are (the 2-D mapping left 27 % of the threads idle at 35 x 48 vectors), and consecutive threads read
  // consecutive 16 bytes of the input row
  const int cv = C >> 3;
  const int flat = blockIdx.y * blockDim.x + threadIdx.x;
  if (flat >= Wo * cv) return;
  const int wo = flat / cv, c8 = flat - wo * cv;
  const int n = blockIdx.x / nbands, b = blockIdx.x - n * nbands;
  const int ho0 = b * band, ho1 = min(ho0 + band, Ho);
  const int xb = wo * S - pad;
  const int ncols = min(xb + K, W) - max(xb, 0);
  const uint4* img = reinterpret_cast<const uint4*>(in + (size_t)n * iHp * iWp * C) + ((size_t)ip * iWp + ip) * cv + c8;
  uint4* obase = reinterpret_cast<uint4*>(out + (((size_t)n * oHp + op) * oWp + op) * out_C + out_c_off) + c8;
  const int ocv = out_C >> 3;
  float bz[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) bz[i] = pbias ? __ldg(pbias + c8 * 8 + i) : 0.f;
  PoolRow<IS_MAX> h[K];
  auto hrow = [&](int y, PoolRow<IS_MAX>& r) {
    r.clear();
    if (y < 0 || y >= H) return;
    const uint4* row = img + (size_t)y * iWp * cv;
#pragma unroll
    for (int dx = 0; dx < K; ++dx) {
      const int x = xb + dx;
      if (x >= 0 && x < W) r.add(__ldg(row + (size_t)x * cv));
    }
  };
  constexpr int KEEP = K > S ? K - S : 0;      // rows shared by consecutive windows
  {
    const int yb = ho0 * S - pad;
#pragma unroll
    for (int i = 0; i < KEEP; ++i) hrow(yb + i, h[i + S]);
  }
#pragma unroll 2
  for (int ho = ho0; ho < ho1; ++ho) {
    const int yb = ho * S - pad;
#pragma unroll
    for (int i = 0; i < KEEP; ++i) h[i] = h[i + S];
#pragma unroll
    for (int i = KEEP; i < K; ++i) hrow(yb + i, h[i]);
    PoolRow<IS_MAX> m = h[0];
#pragma unroll
    for (int i = 1; i < K; ++i) m.merge(h[i]);
    float rcp = 1.f;
    if (!IS_MAX) {
      const int nrows = min(yb + K, H) - max(yb, 0);
      rcp = 1.0f / (count_include_pad ? (float)(K * K) : (float)(nrows * ncols));
    }
    obase[((size_t)ho * oWp + wo) * ocv] = pbias ? m.result_bias(rcp, bz, prelu) : m.result(rcp);
  }
}

// 3x3 pools, input rows staged through shared memory by TMA.  The register-ring kernel above keeps only the loads
// of one row iteration in flight and two of its three loads per row hit lines a neighbour already requested, so the
// unique bytes in flight per SM (~24 KB) cover about half of the HBM latency-bandwidth product (2.7-3.0 TB/s
// measured).  Here a CTA owns a (wt output columns) x (ct 16-byte channel vectors) tile of one image and walks down
// all its rows; one elected thread streams the matching input box of every row ((wt-1)*S+3 columns x ct vectors,
// one 3-D TMA box, out-of-bounds columns zero-filled and masked by the window bounds) into an R-deep ring with
// mbarrier completion: R - 1 rows per CTA are always in flight, tiles span the whole image width where they can
// (no halo columns re-read), and the threads keep the K-row register ring of the kernel above.
template <bool IS_MAX, int S>
__global__ void __launch_bounds__(288)
    k_pool3_tma(const __grid_constant__ CUtensorMap tmI, int H, int W, int pad, int count_include_pad, int Ho, int Wo,
                __nv_bfloat16* __restrict__ out, int out_C, int out_c_off, int iHp, int ip, int oHp, int oWp, int op,
                int ct, int wt, int row_bytes, int row_stride, int R) {
  constexpr int K = 3;
  extern __shared__ __align__(128) uint8_t pool_smem[];
  const uint32_t base = (ptx::smem_u32(pool_smem) + 127u) & ~127u;
  uint8_t* sm = pool_smem + (base - ptx::smem_u32(pool_smem));
  uint64_t* full = reinterpret_cast<uint64_t*>(sm);                  // R <= 16 barriers
  uint8_t* ring = sm + 128;                                          // R rows of row_stride bytes (128-aligned)
  const int n = blockIdx.x, wo0 = blockIdx.y * wt, c0 = blockIdx.z * ct;
  const int wl = (int)threadIdx.x / ct, cl = (int)threadIdx.x - wl * ct;
  const int wo = wo0 + wl;
  const bool active = wl < wt && wo < Wo;
  const int x0 = wo0 * S - pad;                                      // image column of box column 0 (may be -1)
  const int y_begin = max(-pad, 0), y_end = min((Ho - 1) * S - pad + K, H);     // valid input rows [y_begin, y_end)
  const int nrows = y_end - y_begin;
  auto issue = [&](int k) {                   // k-th valid row -> ring slot k % R (thread 0 only)
    const int st = k % R;
    ptx::mbar_arrive_expect_tx(&full[st], (uint32_t)row_bytes);
    ptx::tma_load_3d(ring + (size_t)st * row_stride, &tmI, &full[st], c0 * 8, ip + x0, n * iHp + ip + y_begin + k);
  };
  if (threadIdx.x == 0) {
    ptx::prefetch_tmap(&tmI);
    for (int i = 0; i < R; ++i) ptx::mbar_init(&full[i], 1);
    ptx::fence_barrier_init();
    for (int k = 0; k < R && k < nrows; ++k) issue(k);
  }
  __syncthreads();
  const int xb = wo * S - pad;
  const int ncols = min(xb + K, W) - max(xb, 0);
  uint4* obase = reinterpret_cast<uint4*>(out + (((size_t)n * oHp + op) * oWp + op) * out_C + out_c_off) + c0 + cl;
  const int ocv = out_C >> 3;
  int kk = 0, st = 0;
  uint32_t par = 0;
  PoolRow<IS_MAX> h[K];
  auto hrow = [&](int y, PoolRow<IS_MAX>& r) {     // rows are consumed in increasing y: the next one is ring slot `st`
    r.clear();
    if (y < 0 || y >= H) return;
    ptx::mbar_wait(&full[st], par);
    const uint4* row = reinterpret_cast<const uint4*>(ring + (size_t)st * row_stride) + (wl * S) * ct + cl;
    if (active) {
#pragma unroll
      for (int dx = 0; dx < K; ++dx) {
        const int x = xb + dx;
        if (x >= 0 && x < W) r.add(row[dx * ct]);
      }
    }
    __syncthreads();                               // slot read by everyone: refill it with row kk + R
    if (threadIdx.x == 0 && kk + R < nrows) issue(kk + R);
    ++kk;
    if (++st == R) { st = 0; par ^= 1; }
  };
  constexpr int KEEP = K - S;
  {
    const int yb = -pad;
#pragma unroll
    for (int i = 0; i < KEEP; ++i) hrow(yb + i, h[i + S]);
  }
  for (int ho = 0; ho < Ho; ++ho) {
    const int yb = ho * S - pad;
#pragma unroll
    for (int i = 0; i < KEEP; ++i) h[i] = h[i + S];
#pragma unroll
    for (int i = KEEP; i < K; ++i) hrow(yb + i, h[i]);
    PoolRow<IS_MAX> m = h[0];
#pragma unroll
    for (int i = 1; i < K; ++i) m.merge(h[i]);
    float rcp = 1.f;
    if (!IS_MAX) {
      const int nr = min(yb + K, H) - max(yb, 0);
      rcp = 1.0f / (count_include_pad ? (float)(K * K) : (float)(nr * ncols));
    }
    if (active) obase[((size_t)ho * oWp + wo) * ocv] = m.result(rcp);
  }
}

// Stride-1 3x3 pools (every Inception average pool): the kernel above is ISSUE-bound there (ncu: 90 % issue-slot
// utilisation at 4.1 TB/s, ~220 instructions per output vector: three loads + unpacks per output, register-ring
// moves, one IEEE division).  Same TMA ring, but a thread owns TWO adjacent output columns (four input columns per
// row, the middle pair summed once), the three-row ring is addressed statically (row loop unrolled by 3, no moves)
// and the 1/count factors come from a three-entry per-thread table.
template <bool IS_MAX>
__global__ void __launch_bounds__(288, 2)
    k_pool3s1_tma(const __grid_constant__ CUtensorMap tmI, int H, int W, int pad, int count_include_pad, int Ho, int Wo,
                  __nv_bfloat16* __restrict__ out, int out_C, int out_c_off, int iHp, int ip, int oHp, int oWp, int op,
                  int ct, int wt, int row_bytes, int row_stride, int R, const float* __restrict__ pbias, int prelu) {
  constexpr int K = 3;
  extern __shared__ __align__(128) uint8_t pool_smem[];
  const uint32_t base = (ptx::smem_u32(pool_smem) + 127u) & ~127u;
  uint8_t* sm = pool_smem + (base - ptx::smem_u32(pool_smem));
  uint64_t* full = reinterpret_cast<uint64_t*>(sm);
  uint8_t* ring = sm + 128;
  const int n = blockIdx.x, wo0 = blockIdx.y * wt, c0 = blockIdx.z * ct;
  const int wl = (int)threadIdx.x / ct, cl = (int)threadIdx.x - wl * ct;       // wl: PAIR of output columns
  const int wo_end = min(wo0 + wt, Wo);
  const int wa = wo0 + 2 * wl;                                                 // first output column of the thread
  const bool act_a = wa < wo_end, act_b = wa + 1 < wo_end;
  const int x0 = wo0 - pad;                                                    // image column of box column 0
  const int y_begin = max(-pad, 0), y_end = min(Ho - 1 - pad + K, H);
  const int nrows = y_end - y_begin;
  auto issue = [&](int k) {
    const int st = k % R;
    ptx::mbar_arrive_expect_tx(&full[st], (uint32_t)row_bytes);
    ptx::tma_load_3d(ring + (size_t)st * row_stride, &tmI, &full[st], c0 * 8, ip + x0, n * iHp + ip + y_begin + k);
  };
  if (threadIdx.x == 0) {
    ptx::prefetch_tmap(&tmI);
    for (int i = 0; i < R; ++i) ptx::mbar_init(&full[i], 1);
    ptx::fence_barrier_init();
    for (int k = 0; k < R && k < nrows; ++k) issue(k);
  }
  __syncthreads();
  const int xa = wa - pad;                                                     // leftmost input column of output a
  bool okc[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) okc[i] = act_a && xa + i >= 0 && xa + i < W;
  float rcp_a[4], rcp_b[4];                                                    // index = valid rows of the window
  {
    const int nca = min(xa + K, W) - max(xa, 0), ncb = min(xa + 1 + K, W) - max(xa + 1, 0);
#pragma unroll
    for (int r = 1; r <= 3; ++r) {
      rcp_a[r] = 1.0f / (count_include_pad ? 9.0f : (float)(r * max(nca, 1)));
      rcp_b[r] = 1.0f / (count_include_pad ? 9.0f : (float)(r * max(ncb, 1)));
    }
    rcp_a[0] = rcp_b[0] = 0.f;
  }
  uint4* obase = reinterpret_cast<uint4*>(out + (((size_t)n * oHp + op) * oWp + op) * out_C + out_c_off) + c0 + cl;
  const int ocv = out_C >> 3;
  float bz[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) bz[i] = pbias ? __ldg(pbias + (c0 + cl) * 8 + i) : 0.f;
  int kk = 0, st = 0;
  uint32_t par = 0;
  PoolRow<IS_MAX> ha[K], hb[K];
  // horizontal pass of input row y for both outputs: a = c0+c1+c2, b = c1+c2+c3 (masked columns are neutral)
  auto hrow = [&](int y, PoolRow<IS_MAX>& ra, PoolRow<IS_MAX>& rb) {
    ra.clear();
    rb.clear();
    if (y < 0 || y >= H) return;
    ptx::mbar_wait(&full[st], par);
    const uint4* row = reinterpret_cast<const uint4*>(ring + (size_t)st * row_stride) + (2 * wl) * ct + cl;
    if (act_a) {
      const uint4 nt = PoolRow<IS_MAX>::neutral();
      const uint4 v0 = okc[0] ? row[0] : nt, v1 = okc[1] ? row[1 * ct] : nt;
      const uint4 v2 = okc[2] ? row[2 * ct] : nt, v3 = okc[3] ? row[3 * ct] : nt;
      PoolRow<IS_MAX> mid;
      mid.set(v1);
      mid.add(v2);
      ra.set(v0);
      ra.merge(mid);
      rb.set(v3);
      rb.merge(mid);
    }
    __syncthreads();
    if (threadIdx.x == 0 && kk + R < nrows) issue(kk + R);
    ++kk;
    if (++st == R) { st = 0; par ^= 1; }
  };
  hrow(-pad + 0, ha[0], hb[0]);
  hrow(-pad + 1, ha[1], hb[1]);
  for (int hb0 = 0; hb0 < Ho; hb0 += 3) {
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      const int ho = hb0 + j;
      if (ho < Ho) {                                      // uniform across the CTA
        const int yb = ho - pad;
        hrow(yb + 2, ha[(j + 2) % 3], hb[(j + 2) % 3]);   // input row ho + 2 (relative) lives in ring slot (ho + 2) % 3
        PoolRow<IS_MAX> ma = ha[j % 3], mb = hb[j % 3];
        ma.merge(ha[(j + 1) % 3]);
        ma.merge(ha[(j + 2) % 3]);
        mb.merge(hb[(j + 1) % 3]);
        mb.merge(hb[(j + 2) % 3]);
        const int nr = min(yb + K, H) - max(yb, 0);
        const float fa = nr >= 3 ? rcp_a[3] : (nr == 2 ? rcp_a[2] : rcp_a[1]);
        const float fb = nr >= 3 ? rcp_b[3] : (nr == 2 ? rcp_b[2] : rcp_b[1]);
        uint4* o = obase + ((size_t)ho * oWp + wa) * ocv;
        if (act_a) o[0] = pbias ? ma.result_bias(fa, bz, prelu) : ma.result(fa);
        if (act_b) o[ocv] = pbias ? mb.result_bias(fb, bz, prelu) : mb.result(fb);
      }
    }
  }
}

// K4d: global average pool + FC + softmax (K>1) / sigmoid (K==1); one CTA per image.
// run_inference.py:548-551 does the softmax on the CPU; here it is fused into the tail.
constexpr int HEAD_THREADS = 256;
constexpr int HEAD_MAXK = 16;
__global__ void __launch_bounds__(HEAD_THREADS)
    k_head(const __nv_bfloat16* __restrict__ in, int HW, int C, const float* __restrict__ fcw,
           const float* __restrict__ fcb, int K, float* __restrict__ probs) {
  extern __shared__ float s_feat[];   // C pooled features
  __shared__ float s_red[HEAD_MAXK][HEAD_THREADS / 32];
  __shared__ float s_logit[HEAD_MAXK];
  const int n = blockIdx.x;
  const __nv_bfloat16* x = in + (size_t)n * HW * C;
  for (int c = threadIdx.x; c < C; c += HEAD_THREADS) {
    float s = 0.f;
    for (int p = 0; p < HW; ++p) s += __bfloat162float(x[(size_t)p * C + c]);
    s_feat[c] = s / (float)HW;
  }
  __syncthreads();
  for (int k = 0; k < K; ++k) {
    float s = 0.f;
    for (int c = threadIdx.x; c < C; c += HEAD_THREADS) s += s_feat[c] * fcw[(size_t)k * C + c];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) s_red[k][threadIdx.x >> 5] = s;
  }
  __syncthreads();
  if (threadIdx.x < K) {
    float s = fcb[threadIdx.x];
    for (int w = 0; w < HEAD_THREADS / 32; ++w) s += s_red[threadIdx.x][w];
    s_logit[threadIdx.x] = s;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    if (K == 1) {
      probs[n] = 1.0f / (1.0f + expf(-s_logit[0]));
    } else {
      float mx = s_logit[0];
      for (int k = 1; k < K; ++k) mx = fmaxf(mx, s_logit[k]);
      float e[HEAD_MAXK], sum = 0.f;
      for (int k = 0; k < K; ++k) {
        e[k] = expf(s_logit[k] - mx);
        sum += e[k];
      }
      for (int k = 0; k < K; ++k) probs[(size_t)n * K + k] = e[k] / sum;
    }
  }
}

// dense [B][H][W][C] <-> interior of a padded [B][H+2p][W+2p][C] tensor, 16 bytes per thread
__global__ void k_copy_interior(const uint4* __restrict__ src, uint4* __restrict__ dst, int B, int H, int W, int C8,
                                int Hp, int Wp, int pad, int to_padded) {
  size_t total = (size_t)B * H * W * C8;
  size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  int c = idx % C8;
  size_t p = idx / C8;
  int w = p % W;
  p /= W;
  int h = p % H;
  int n = p / H;
  size_t pi = (((size_t)n * Hp + h + pad) * Wp + w + pad) * C8 + c;
  if (to_padded) dst[pi] = src[idx]; else dst[idx] = src[pi];
}

// Pixel box (bw x bh x bn images == 128 GEMM rows) of one M tile: the shape that wastes the fewest rows
// on partial tiles for this output size and batch; ties go to the wider box (longer contiguous runs).
void pick_box(int Wo, int Ho, int B, int* bw, int* bh, int* bn) {
  double best = 1e30;
  for (int w = 16; w >= 1; w >>= 1)
    for (int h = 16; h >= 1; h >>= 1) {
      if (w * h > 128) continue;
      const int n = 128 / (w * h);
      const double rows = (double)((Wo + w - 1) / w * w) * ((Ho + h - 1) / h * h) * ((B + n - 1) / n * n);
      if (rows < best * 0.999) {
        best = rows;
        *bw = w; *bh = h; *bn = n;
      }
    }
}

int pick_ntile(int Cout) {
  // largest multiple of 32 that is <= 256 and divides Cout
  for (int n = 256; n >= 32; n -= 32)
    if (Cout % n == 0) return n;
  return 0;
}

}  // namespace

extern "C" {

int wsi_net_create(int max_batch, void** net_out) {
  WSI_ARG_CHECK(net_out && max_batch > 0, "wsi_net_create: bad arguments");
  Net* n = new Net();
  n->max_batch = max_batch;
  int dev = 0;
  cudaGetDevice(&dev);
  int sms = 0;
  if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && sms > 0) n->num_sms = sms;
  const char* nf = getenv("WSI_B200_NO_FLAT");
  n->flat_enabled = !(nf && nf[0] == '1');
  const char* nc = getenv("WSI_B200_NO_CLUSTER");
  n->cluster_enabled = !(nc && nc[0] == '1');
  const char* nfp = getenv("WSI_B200_NO_FLAT_PAIR");
  n->no_flat_pair = nfp && nfp[0] == '1';
  n->res_smem_enabled = getenv("WSI_B200_NO_RES_SMEM") == nullptr;
  const char* npr = getenv("WSI_B200_NO_POOL_RING");
  n->pool_ring = !(npr && npr[0] == '1');
  const char* le = getenv("WSI_B200_LEGACY_EPI");
  n->legacy_epi = le && le[0] == '1';
  const char* nsp = getenv("WSI_B200_NO_STEM_POOL");
  n->stem_pool_enabled = !(nsp && nsp[0] == '1');
  const char* npdl = getenv("WSI_B200_NO_PDL");
  n->pdl_enabled = !(npdl && npdl[0] == '1');
  const char* k3 = getenv("WSI_B200_KW3");
  n->kw3_enabled = k3 && k3[0] == '1';
  *net_out = n;
  return WSI_OK;
}

int wsi_net_add_tensor(void* net, int H, int W, int C, int external, int* id_out) {
  Net* n = (Net*)net;
  WSI_ARG_CHECK(n && id_out && !n->finalized, "wsi_net_add_tensor: bad state");
  WSI_ARG_CHECK(H > 0 && W > 0 && C > 0, "wsi_net_add_tensor: bad shape");
  Tensor t;
  t.H = H; t.W = W; t.C = C; t.external = external != 0;
  n->tensors.push_back(t);
  *id_out = (int)n->tensors.size() - 1;
  return WSI_OK;
}

// generic conv: w is host fp32 OIHW [Cout][Cin][KH][KW]; bias host fp32 [Cout] or NULL
int wsi_net_add_conv(void* net, int in_id, int out_id, int out_c_off, const float* w_oihw, const float* bias,
                     int Cin, int Cout, int KH, int KW, int stride, int pad_h, int pad_w, int relu,
                     int residual_id) {
  return wsi_net_add_conv_slice(net, in_id, 0, out_id, out_c_off, w_oihw, bias, Cin, Cout, KH, KW, stride, pad_h, pad_w,
                                relu, residual_id);
}

// Convolution over channels [in_c_off, in_c_off + Cin) of the input tensor (sibling 1x1 convs merged into one launch
// write one wide tensor; their consumers read slices of it).  The A-operand tensor maps start at the slice and
// declare Cin channels, so TMA's zero fill past the slice works as for a whole tensor.
int wsi_net_add_conv_slice(void* net, int in_id, int in_c_off, int out_id, int out_c_off, const float* w_oihw,
                           const float* bias, int Cin, int Cout, int KH, int KW, int stride, int pad_h, int pad_w,
                           int relu, int residual_id) {
  Net* n = (Net*)net;
  WSI_ARG_CHECK(n && !n->finalized && w_oihw, "wsi_net_add_conv: bad state");
  WSI_ARG_CHECK(in_id >= 0 && in_id < (int)n->tensors.size() && out_id >= 0 && out_id < (int)n->tensors.size(),
                "wsi_net_add_conv: bad tensor id");
  const Tensor& ti = n->tensors[in_id];
  const Tensor& to = n->tensors[out_id];
  WSI_ARG_CHECK(in_c_off >= 0 && in_c_off % 8 == 0 && in_c_off + Cin <= ti.C && (in_c_off == 0 || !ti.external),
                "wsi_net_add_conv: channels [%d, %d) outside tensor C %d", in_c_off, in_c_off + Cin, ti.C);
  WSI_ARG_CHECK(ti.C == Cin || !ti.external, "wsi_net_add_conv: Cin %d != external tensor C %d", Cin, ti.C);
  WSI_ARG_CHECK(Cin % 8 == 0, "wsi_net_add_conv: Cin must be a multiple of 8 (got %d)", Cin);
  const int Ho = (ti.H + 2 * pad_h - KH) / stride + 1, Wo = (ti.W + 2 * pad_w - KW) / stride + 1;
  WSI_ARG_CHECK(Ho == to.H && Wo == to.W, "wsi_net_add_conv: output shape %dx%d != tensor %dx%d", Ho, Wo, to.H,
                to.W);
  WSI_ARG_CHECK(out_c_off >= 0 && out_c_off + Cout <= to.C && out_c_off % 8 == 0 && to.C % 8 == 0,
                "wsi_net_add_conv: bad channel offset");
  WSI_ARG_CHECK(pick_ntile(Cout) > 0, "wsi_net_add_conv: Cout %d must be a multiple of 32", Cout);
  if (residual_id >= 0) {
    WSI_ARG_CHECK(residual_id < (int)n->tensors.size(), "wsi_net_add_conv: bad residual id");
    const Tensor& tr = n->tensors[residual_id];
    WSI_ARG_CHECK(tr.H == to.H && tr.W == to.W && tr.C == Cout && out_c_off == 0,
                  "wsi_net_add_conv: residual shape mismatch");
  }
  Op op;
  op.kind = OP_CONV;
  op.in = in_id; op.out = out_id; op.res = residual_id;
  op.Cin = Cin; op.Cout = Cout; op.KH = KH; op.KW = KW; op.stride = stride; op.pad_h = pad_h; op.pad_w = pad_w;
  op.relu = relu; op.out_c_off = out_c_off; op.is_stem = 0; op.in_c_off = in_c_off;
  const int chunks = (Cin + 63) / 64;
  op.Ktot = KH * KW * chunks * 64;
  op.h_w.assign((size_t)Cout * op.Ktot, __float2bfloat16(0.f));
  for (int o = 0; o < Cout; ++o)
    for (int c = 0; c < Cin; ++c)
      for (int y = 0; y < KH; ++y)
        for (int x = 0; x < KW; ++x) {
          float v = w_oihw[(((size_t)o * Cin + c) * KH + y) * KW + x];
          size_t k = ((size_t)(y * KW + x) * chunks + c / 64) * 64 + (c % 64);
          op.h_w[(size_t)o * op.Ktot + k] = __float2bfloat16(v);
        }
  op.cin_real = 0;
  for (int c = Cin - 1; c >= 0 && !op.cin_real; --c)
    for (int o = 0; o < Cout && !op.cin_real; ++o)
      for (int t = 0; t < KH * KW; ++t)
        if (w_oihw[((size_t)o * Cin + c) * KH * KW + t] != 0.f) {
          op.cin_real = c + 1;
          break;
        }
  op.h_bias.assign(Cout, 0.f);
  if (bias) memcpy(op.h_bias.data(), bias, sizeof(float) * Cout);
  op.flops_per_image = 2.0 * Ho * Wo * (double)Cout * Cin * KH * KW;
  n->ops.push_back(std::move(op));
  return WSI_OK;
}

// First convolution on the 3-channel image through the parity-plane stem kernel (conv_stem.cuh):
//   ksize 7, pad 3 (every ResNet): input tensor = the preprocess kernel's padded layout [B][S+6][S+6][4];
//   ksize 3, pad 0 (InceptionV4 conv2d_1a): input tensor [B][Se][Se][4], Se = S rounded up to even.
// Stride is 2.  Cout == 64 (or 32 for ksize 3) takes the parity kernel; a 7x7 stem of another width falls back to
// the per-tap kernel's overlapping-window map.
int wsi_net_add_stem_parity(void* net, int in_id, int out_id, const float* w_oihw, const float* bias, int Cout,
                            int ksize, int pad, int relu) {
  Net* n = (Net*)net;
  WSI_ARG_CHECK(n && !n->finalized && w_oihw, "wsi_net_add_stem_parity: bad state");
  WSI_ARG_CHECK(in_id >= 0 && in_id < (int)n->tensors.size() && out_id >= 0 && out_id < (int)n->tensors.size(),
                "wsi_net_add_stem_parity: bad tensor id");
  WSI_ARG_CHECK((ksize == 7 && pad == 3) || (ksize == 3 && pad == 0), "wsi_net_add_stem_parity: 7x7/pad 3 or 3x3/pad 0");
  const Tensor& ti = n->tensors[in_id];
  const Tensor& to = n->tensors[out_id];
  WSI_ARG_CHECK(ti.C == 4 && (ti.W % 2) == 0 && (ti.H % 2) == 0 && ti.H == ti.W,
                "wsi_net_add_stem_parity: input must be square NHWC4 with an even side");
  const int Ho = (ti.H - ksize) / 2 + 1;
  WSI_ARG_CHECK(to.H == Ho && to.W == Ho && to.C == Cout, "wsi_net_add_stem_parity: bad output tensor");
  WSI_ARG_CHECK(pick_ntile(Cout) > 0, "wsi_net_add_stem_parity: Cout must be a multiple of 32");
  const bool parity = n->flat_enabled && (Cout == 64 || (ksize == 3 && Cout == 32));
  WSI_ARG_CHECK(parity || ksize == 7, "wsi_net_add_stem_parity: no kernel for a 3x3 stem with %d filters", Cout);
  Op op;
  op.kind = OP_CONV;
  op.in = in_id; op.out = out_id; op.res = -1;
  op.Cin = 3; op.Cout = Cout; op.KH = ksize; op.KW = ksize; op.stride = 2; op.pad_h = pad; op.pad_w = pad;
  op.relu = relu; op.out_c_off = 0;
  op.is_stem = parity ? 2 : 1;
  const int JJ = ksize == 7 ? 2 : 1;
  op.Ktot = parity ? ksize * JJ * 16 : 7 * 32;
  op.h_w.assign((size_t)Cout * op.Ktot, __float2bfloat16(0.f));
  for (int o = 0; o < Cout; ++o)
    for (int c = 0; c < 3; ++c)
      for (int y = 0; y < ksize; ++y)
        for (int x = 0; x < ksize; ++x) {
          float v = w_oihw[(((size_t)o * 3 + c) * ksize + y) * ksize + x];
          size_t off;
          if (op.is_stem == 2) {
            // canonical no-swizzle blocks [ks = JJ*kh + jj][n/8][k/8][n%8][k%8], k = 16 per block:
            // pixel pair j = x/2 = 2*jj + k/8, element (x%2)*4 + c inside the pair
            const int j = x >> 1, ks = y * JJ + (j >> 1), k16 = (j & 1) * 8 + (x & 1) * 4 + c;
            off = (size_t)ks * Cout * 16 + (size_t)(o / 8) * 128 + (size_t)(k16 / 8) * 64 + (size_t)(o % 8) * 8 + (k16 % 8);
          } else {
            // K layout: chunk kh x [px 0..7][ch 0..3]; px==7 and ch==3 carry zero weights
            off = (size_t)o * op.Ktot + (size_t)y * 32 + x * 4 + c;
          }
          op.h_w[off] = __float2bfloat16(v);
        }
  op.h_bias.assign(Cout, 0.f);
  if (bias) memcpy(op.h_bias.data(), bias, sizeof(float) * Cout);
  op.flops_per_image = 2.0 * to.H * to.W * (double)Cout * 3 * ksize * ksize;
  n->ops.push_back(std::move(op));
  return WSI_OK;
}

int wsi_net_add_stem7x7(void* net, int in_id, int out_id, const float* w_oihw, const float* bias, int Cout,
                        int relu) {
  return wsi_net_add_stem_parity(net, in_id, out_id, w_oihw, bias, Cout, 7, 3, relu);
}

int wsi_net_add_pool(void* net, int in_id, int out_id, int out_c_off, int is_max, int k, int stride, int pad,
                     int count_include_pad) {
  Net* n = (Net*)net;
  WSI_ARG_CHECK(n && !n->finalized, "wsi_net_add_pool: bad state");
  WSI_ARG_CHECK(in_id >= 0 && in_id < (int)n->tensors.size() && out_id >= 0 && out_id < (int)n->tensors.size(),
                "wsi_net_add_pool: bad tensor id");
  const Tensor& ti = n->tensors[in_id];
  const Tensor& to = n->tensors[out_id];
  const int Ho = (ti.H + 2 * pad - k) / stride + 1, Wo = (ti.W + 2 * pad - k) / stride + 1;
  WSI_ARG_CHECK(Ho == to.H && Wo == to.W, "wsi_net_add_pool: output shape mismatch");
  WSI_ARG_CHECK(ti.C % 8 == 0 && to.C % 8 == 0 && out_c_off % 8 == 0 && out_c_off + ti.C <= to.C,
                "wsi_net_add_pool: channel layout");
  Op op;
  op.kind = is_max ? OP_MAXPOOL : OP_AVGPOOL;
  op.in = in_id; op.out = out_id; op.res = -1;
  op.pk = k; op.pstride = stride; op.ppad = pad; op.count_include_pad = count_include_pad;
  op.out_c_off = out_c_off;
  n->ops.push_back(std::move(op));
  return WSI_OK;
}

// ReLU of the conv added last only on its first `cols` output channels (multiple of 64).
int wsi_net_set_conv_relu_cols(void* net, int cols) {
  Net* n = (Net*)net;
  WSI_ARG_CHECK(n && !n->finalized && !n->ops.empty() && n->ops.back().kind == OP_CONV, "wsi_net_set_conv_relu_cols: no conv");
  Op& op = n->ops.back();
  WSI_ARG_CHECK(op.relu && cols > 0 && cols <= op.Cout && cols % 64 == 0,
                "wsi_net_set_conv_relu_cols: bad column count %d", cols);
  op.relu_cols = cols;
  return WSI_OK;
}

// Residual of the conv added last = channels [res_c_off, res_c_off + Cout) of tensor `residual_id` (the conv must have
// been added without a residual).  Lets a BasicBlock's second conv add the downsample branch where it lies inside the
// tensor it shares with the block's first conv (wsi_net_set_conv_center_cols).
int wsi_net_set_conv_residual_slice(void* net, int residual_id, int res_c_off) {
  Net* n = (Net*)net;
  WSI_ARG_CHECK(n && !n->finalized && !n->ops.empty() && n->ops.back().kind == OP_CONV, "wsi_net_set_conv_residual_slice: no conv");
  Op& op = n->ops.back();
  WSI_ARG_CHECK(op.res < 0 && !op.is_stem && op.out_c_off == 0, "wsi_net_set_conv_residual_slice: conv already has a residual");
  WSI_ARG_CHECK(residual_id >= 0 && residual_id < (int)n->tensors.size(), "wsi_net_set_conv_residual_slice: bad residual id");
  const Tensor& tr = n->tensors[residual_id];
  const Tensor& to = n->tensors[op.out];
  WSI_ARG_CHECK(tr.H == to.H && tr.W == to.W && !tr.external && res_c_off >= 0 && res_c_off % 8 == 0 &&
                    res_c_off + op.Cout <= tr.C,
                "wsi_net_set_conv_residual_slice: channels [%d, %d) of a %dx%dx%d tensor do not match the %dx%dx%d output",
                res_c_off, res_c_off + op.Cout, tr.H, tr.W, tr.C, to.H, to.W, op.Cout);
  op.res = residual_id;
  op.res_c_off = res_c_off;
  return WSI_OK;
}

// Output channels >= cols of the conv added last have non-zero weights at the CENTRE tap only (a 1x1 / stride-s / pad-0
// convolution embedded in a KxK / stride-s / pad-(K-1)/2 one that reads the same pixels: the downsample branch of a
// ResNet BasicBlock next to the block's first conv).  The per-tap kernel gives those channels their own N tiles and
// runs only the centre tap's k-blocks for them, so the block input is read by ONE launch.
int wsi_net_set_conv_center_cols(void* net, int cols) {
  Net* n = (Net*)net;
  WSI_ARG_CHECK(n && !n->finalized && !n->ops.empty() && n->ops.back().kind == OP_CONV, "wsi_net_set_conv_center_cols: no conv");
  Op& op = n->ops.back();
  WSI_ARG_CHECK(!op.is_stem && (op.KH & 1) && (op.KW & 1) && op.pad_h == op.KH / 2 && op.pad_w == op.KW / 2,
                "wsi_net_set_conv_center_cols: needs an odd kernel with 'same' padding");
  WSI_ARG_CHECK(cols > 0 && cols < op.Cout && cols % 64 == 0 && (op.Cout - cols) % 64 == 0,
                "wsi_net_set_conv_center_cols: bad column count %d of %d", cols, op.Cout);
  const int chunks = (op.Cin + 63) / 64, ctap = op.pad_h * op.KW + op.pad_w;
  for (int o = cols; o < op.Cout; ++o)
    for (int t = 0; t < op.KH * op.KW; ++t) {
      if (t == ctap) continue;
      for (int k = 0; k < chunks * 64; ++k)
        WSI_ARG_CHECK(__bfloat162float(op.h_w[(size_t)o * op.Ktot + (size_t)t * chunks * 64 + k]) == 0.f,
                      "wsi_net_set_conv_center_cols: channel %d has a weight outside the centre tap", o);
    }
  op.center_cols = cols;
  const Tensor& to = n->tensors[op.out];
  op.flops_per_image = 2.0 * to.H * to.W * (double)op.Cin * ((double)cols * op.KH * op.KW + (double)(op.Cout - cols));
  return WSI_OK;
}

// 3x3 / stride 1 AVERAGE pool followed by + bias[c] and optional ReLU.  conv1x1(avgpool(x)) == avgpool(conv1x1(x))
// (both linear, the divisor depends on the position only), so the engine runs the 1x1 conv of an Inception pool branch
// FIRST (without bias / ReLU) and pools its 4-8x narrower output here: the wide block input is no longer read and
// re-written by a pooling pass.
int wsi_net_add_pool_bias(void* net, int in_id, int in_c_off, int C, int out_id, int out_c_off, int k, int stride, int pad,
                          int count_include_pad, const float* bias, int relu) {
  Net* n = (Net*)net;
  WSI_ARG_CHECK(n && !n->finalized && k == 3 && stride == 1 && bias,
                "wsi_net_add_pool_bias: only 3x3 / stride 1 average pools");
  WSI_ARG_CHECK(in_id >= 0 && in_id < (int)n->tensors.size() && out_id >= 0 && out_id < (int)n->tensors.size(),
                "wsi_net_add_pool_bias: bad tensor id");
  const Tensor& ti = n->tensors[in_id];
  const Tensor& to = n->tensors[out_id];
  const int Ho = ti.H + 2 * pad - k + 1, Wo = ti.W + 2 * pad - k + 1;
  WSI_ARG_CHECK(Ho == to.H && Wo == to.W, "wsi_net_add_pool_bias: output shape mismatch");
  WSI_ARG_CHECK(C > 0 && C % 8 == 0 && in_c_off >= 0 && in_c_off % 8 == 0 && in_c_off + C <= ti.C && !ti.external &&
                    to.C % 8 == 0 && out_c_off % 8 == 0 && out_c_off + C <= to.C,
                "wsi_net_add_pool_bias: channel layout");
  Op op;
  op.kind = OP_AVGPOOL;
  op.in = in_id; op.out = out_id; op.res = -1;
  op.pk = k; op.pstride = stride; op.ppad = pad; op.count_include_pad = count_include_pad;
  op.out_c_off = out_c_off;
  op.in_c_off = in_c_off;
  op.pC = C;
  op.h_pbias.assign(bias, bias + C);
  op.prelu = relu ? 1 : 0;
  n->ops.push_back(std::move(op));
  return WSI_OK;
}

int wsi_net_add_head(void* net, int in_id, const float* fc_w, const float* fc_b, int K) {
  Net* n = (Net*)net;
  WSI_ARG_CHECK(n && !n->finalized && fc_w, "wsi_net_add_head: bad state");
  WSI_ARG_CHECK(in_id >= 0 && in_id < (int)n->tensors.size(), "wsi_net_add_head: bad tensor id");
  WSI_ARG_CHECK(K >= 1 && K <= HEAD_MAXK, "wsi_net_add_head: K out of range");
  const Tensor& ti = n->tensors[in_id];
  Op op;
  op.kind = OP_HEAD;
  op.in = in_id; op.out = -1; op.res = -1;
  op.K = K;
  op.h_fcw.assign(fc_w, fc_w + (size_t)K * ti.C);
  op.h_fcb.assign(K, 0.f);
  if (fc_b) memcpy(op.h_fcb.data(), fc_b, sizeof(float) * K);
  op.flops_per_image = 2.0 * K * ti.C;
  n->ops.push_back(std::move(op));
  return WSI_OK;
}

// input tensor map of the TMA-ring pool kernel: (C, Wp, B*Hp) bf16, box = (ct vectors, box columns, one row)
static int encode_pool_map(Net* n, Op& op, const void* base) {
  const Tensor& ti = n->tensors[op.in];
  uint64_t dims[3] = {(uint64_t)(op.pC ? op.pC : ti.C), (uint64_t)ti.Wp(), (uint64_t)n->max_batch * ti.Hp()};
  uint64_t strides[2] = {(uint64_t)ti.C * 2, (uint64_t)ti.Wp() * ti.C * 2};
  uint32_t box[3] = {(uint32_t)op.pt_ct * 8, (uint32_t)op.pt_cols, 1};
  uint32_t es[3] = {1, 1, 1};
  base = (const __nv_bfloat16*)base + op.in_c_off;
  return wsi_make_tmap(&op.tmA, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, base, dims, strides, box, es,
                       CU_TENSOR_MAP_SWIZZLE_NONE);
}

int wsi_net_finalize(void* net) {
  Net* n = (Net*)net;
  WSI_ARG_CHECK(n && !n->finalized, "wsi_net_finalize: bad state");
  const int B = n->max_batch;
  for (auto& op : n->ops)
    if (op.kind == OP_CONV && op.is_stem == 2) {
      // the parity-plane stem writes every column / row of its P x P flat grid: P - S/2 junk ones
      Tensor& to = n->tensors[op.out];
      const int P = n->tensors[op.in].H / 2;
      to.xh = P - to.H;
      to.xw = P - to.W;
      to.lock_geom = true;
    }
  // Which convs run on the flat-slab kernel, and the stored geometry that follows from it.  The kernel computes a
  // CENTRED 3x3 window at every position of a stored [Hp][Wp] grid shared by input, output and residual:
  //   pad 1 ('same'):  the input needs a zero border >= 1 and the output keeps the input's border;
  //   pad 0 ('valid'): no border needed and the (two pixels smaller) output sits one pixel further in,
  //                    i.e. its border is the input's + 1 -- which is exactly the zero padding a following
  //                    'same' conv wants.
  // Borders only grow, so the constraints are propagated to a fixed point; an op whose tensors cannot agree
  // (a geometry-locked producer, two producers of one concat buffer that disagree) falls back to the per-tap kernel.
  for (auto& op : n->ops) {
    op.flat = false;
    if (op.kind != OP_CONV || op.is_stem || !n->flat_enabled) continue;
    const Tensor& ti = n->tensors[op.in];
    bool ok = op.KH == 3 && op.KW == 3 && op.stride == 1 && op.pad_h == op.pad_w && (op.pad_h == 0 || op.pad_h == 1) &&
              op.Cin % 8 == 0 && op.Cout % 32 == 0 && op.Cout <= 128 && !ti.external && ti.W >= 24 &&
              ti.H >= 8;
    if (op.pad_h == 0 && getenv("WSI_B200_NO_FLAT_VALID")) ok = false;
    if (op.res >= 0 && n->tensors[op.res].external) ok = false;
    if (op.center_cols >= 0) ok = false;
    op.flat = ok;
  }
  for (int round = 0; round < 64; ++round) {
    for (auto& t : n->tensors)
      if (!t.lock_geom) t.pad = 0;
    bool conflict = false;
    for (int iter = 0; iter < 64 && !conflict; ++iter) {
      bool changed = false;
      for (auto& op : n->ops) {
        if (!op.flat) continue;
        Tensor& ti = n->tensors[op.in];
        Tensor& to = n->tensors[op.out];
        const int p = op.pad_h;
        auto raise = [&](Tensor& t, int want) {
          if (t.pad >= want) return true;
          if (t.lock_geom) return false;
          t.pad = want;
          changed = true;
          return true;
        };
        bool ok = raise(ti, p) && raise(to, ti.pad + 1 - p) && raise(ti, to.pad - 1 + p);
        if (ok && op.res >= 0) {
          Tensor& tr = n->tensors[op.res];
          ok = raise(tr, to.pad) && raise(to, tr.pad) && raise(ti, to.pad - 1 + p);
        }
        if (!ok) {
          op.flat = false;
          conflict = true;
          break;
        }
      }
      if (!changed) break;
    }
    if (conflict) continue;
    // junk rows / columns after the image so that every flat op's tensors share one stored grid
    for (auto& t : n->tensors)
      if (!t.lock_geom) t.xh = t.xw = 0;
    for (auto& op : n->ops) {
      if (!op.flat) continue;
      const Tensor& ti = n->tensors[op.in];
      Tensor& to = n->tensors[op.out];
      const int xh = ti.Hp() - to.H - 2 * to.pad, xw = ti.Wp() - to.W - 2 * to.pad;
      bool ok = xh >= 0 && xw >= 0;
      if (ok && (to.Hp() != ti.Hp() || to.Wp() != ti.Wp())) {
        if (to.lock_geom || to.xh != 0 || to.xw != 0) ok = false;   // already fixed by another producer
        else { to.xh = xh; to.xw = xw; }
      }
      if (ok && op.res >= 0) {
        const Tensor& tr = n->tensors[op.res];
        ok = tr.Hp() == to.Hp() && tr.Wp() == to.Wp();
      }
      if (!ok) {
        op.flat = false;
        conflict = true;
        break;
      }
    }
    if (!conflict) break;
  }
  // Cout = 64 flat convs over short rows (ResNet layer1): the kw-split kernel, whose slab (128 + 2*Wp rows) is one TMA box
  for (auto& op : n->ops) {
    op.flat3 = false;
    if (!op.flat || !n->kw3_enabled || op.Cout != convf3::NC || op.Cin % 64 != 0) continue;
    const Tensor& ti = n->tensors[op.in];
    const int cl = (n->cluster_enabled && !n->no_flat_pair) ? 2 : 1;
    const int rows = ((128 + 2 * ti.Wp() + 7) / 8) * 8;
    if (rows > 256) continue;
    const long long room = 227 * 1024 - (long long)convf3::smem_fixed(op.Cin / 64, cl);
    if (room / ((long long)rows * 128) < 2) continue;
    op.flat3 = true;
  }
  // stem + ReLU + 3x3/2/pad-1 max pooling (every ResNet): fuse when the stem output has no other reader
  if (n->stem_pool_enabled)
    for (size_t i = 0; i < n->ops.size(); ++i) {
      Op& op = n->ops[i];
      if (op.kind != OP_CONV || op.is_stem != 2 || !op.relu || op.KH != 7 || op.Cout != 64) continue;
      int readers = 0, pool_idx = -1;
      for (size_t j = 0; j < n->ops.size(); ++j) {
        const Op& o2 = n->ops[j];
        if (o2.in == op.out || (o2.kind == OP_CONV && o2.res == op.out)) {
          ++readers;
          if (o2.kind == OP_MAXPOOL && o2.pk == 3 && o2.pstride == 2 && o2.ppad == 1 && j > i) pool_idx = (int)j;
        }
      }
      const Tensor& ts = n->tensors[op.out];
      const int P = n->tensors[op.in].H / 2;
      if (readers == 1 && pool_idx >= 0 && ts.H % 2 == 0 && ts.W % 2 == 0 && P <= 128 && (ts.W / 2) * 8 <= 512 &&
          n->tensors[n->ops[pool_idx].out].C % 8 == 0) {
        op.fused_pool = pool_idx;
        n->ops[pool_idx].skip = true;
        n->tensors[op.out].elided = true;
      }
    }
  for (auto& op : n->ops)
    if (op.kind == OP_HEAD && (n->tensors[op.in].pad || n->tensors[op.in].xh || n->tensors[op.in].xw)) {
      wsi_set_error("wsi_net_finalize: the head cannot read a padded tensor");
      return WSI_ERR_UNSUPPORTED;
    }
  for (auto& t : n->tensors) {
    if (t.external || t.elided) continue;
    size_t bytes = (size_t)B * t.Hp() * t.Wp() * t.C * sizeof(__nv_bfloat16);
    WSI_CUDA_CHECK(cudaMalloc(&t.ptr, bytes));
    WSI_CUDA_CHECK(cudaMemset(t.ptr, 0, bytes));
  }
  int max_smem = 0, max_smem_flat = 0, max_smem_flat3 = 0, max_smem_stem = 0, max_smem_stem_pool = 0;
  for (auto& op : n->ops) {
    if (op.kind == OP_CONV) {
      const Tensor& ti = n->tensors[op.in];
      const Tensor& to = n->tensors[op.out];
      if (op.flat3) {
        // kw-split weight matrix [kw*64 + co][(kh*chunks + c)*64 + ci] from the tap-major [co][((kh*3 + kw)*chunks + c)*64 + ci]
        const int chunks = op.Cin / 64, K3 = 3 * chunks * 64;
        std::vector<__nv_bfloat16> w3((size_t)convf3::N3 * K3);
        for (int kw = 0; kw < 3; ++kw)
          for (int co = 0; co < convf3::NC; ++co)
            for (int kh = 0; kh < 3; ++kh)
              for (int c = 0; c < chunks * 64; ++c)
                w3[(size_t)(kw * convf3::NC + co) * K3 + (size_t)(kh * chunks + c / 64) * 64 + c % 64] =
                    op.h_w[(size_t)co * op.Ktot + (size_t)((kh * 3 + kw) * chunks + c / 64) * 64 + c % 64];
        op.h_w.swap(w3);
      }
      WSI_CUDA_CHECK(cudaMalloc(&op.d_w, op.h_w.size() * sizeof(__nv_bfloat16)));
      WSI_CUDA_CHECK(cudaMemcpy(op.d_w, op.h_w.data(), op.h_w.size() * sizeof(__nv_bfloat16),
                                cudaMemcpyHostToDevice));
      WSI_CUDA_CHECK(cudaMalloc(&op.d_bias, op.h_bias.size() * sizeof(float)));
      WSI_CUDA_CHECK(cudaMemcpy(op.d_bias, op.h_bias.data(), op.h_bias.size() * sizeof(float),
                                cudaMemcpyHostToDevice));
      if (op.is_stem == 2) {
        const int P = ti.H / 2;
        convs::Args& sa = op.sa;
        memset(&sa, 0, sizeof(sa));
        const int JJ = op.KH == 7 ? 2 : 1;
        const int w_bytes = op.KH * JJ * op.Cout * 32;
        sa.P = P;
        sa.N = op.Cout;
        sa.plane_pairs = (long long)B * P * P;
        sa.slab_pairs = (convs::T * 128 + ((op.KH - 1) / 2) * P + 2 * JJ + 7) / 8 * 8;
        sa.stages = 4;
        sa.relu = op.relu;
        sa.w = (const uint8_t*)op.d_w;
        sa.bias = op.d_bias;
        if (op.fused_pool >= 0) {
          const Tensor& tp = n->tensors[n->ops[op.fused_pool].out];
          sa.stages = 4;
          sa.row_pairs = to.H / 2;
          sa.pool_w = to.W / 2;
          sa.pool_out = tp.ptr;
          sa.pool_Hp = tp.Hp(); sa.pool_Wp = tp.Wp(); sa.pool_pad = tp.pad; sa.pool_C = tp.C;
          sa.pool_c_off = n->ops[op.fused_pool].out_c_off;
          int smem = (int)convs::smem_bytes(sa.slab_pairs, sa.stages, true, w_bytes);
          WSI_ARG_CHECK(smem <= 227 * 1024, "wsi_net_finalize: fused stem does not fit shared memory (P = %d)", P);
          if (smem > max_smem_stem_pool) max_smem_stem_pool = smem;
          continue;
        }
        int smem = (int)convs::smem_bytes(sa.slab_pairs, sa.stages, false, w_bytes);
        WSI_ARG_CHECK(smem <= 227 * 1024, "wsi_net_finalize: stem slab does not fit shared memory (P = %d)", P);
        if (smem > max_smem_stem) max_smem_stem = smem;
        uint64_t dims[2] = {(uint64_t)to.C, (uint64_t)B * to.Hp() * to.Wp()};
        uint64_t strides[1] = {(uint64_t)to.C * 2};
        uint32_t box[2] = {(uint32_t)op.Cout, 128};
        uint32_t es[2] = {1, 1};
        int rc = wsi_make_tmap(&op.tmO, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, to.ptr, dims, strides, box, es,
                               op.Cout == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B);
        if (rc) return rc;
        continue;
      }
      if (op.flat3) {
        convf3::Args& f = op.f3;
        memset(&f, 0, sizeof(f));
        f.Hp = ti.Hp(); f.Wp = ti.Wp(); f.H = to.H; f.W = to.W;
        f.pad = to.pad;
        f.chunks = op.Cin / 64;
        f.ksteps = (f.chunks == 1 && op.cin_real <= 32) ? 2 : 4;
        f.cluster = (n->cluster_enabled && !n->no_flat_pair) ? 2 : 1;
        f.slab_rows = ((128 + 2 * f.Wp + 7) / 8) * 8;
        f.slab_loads = 1;
        f.slab_box_rows = f.slab_rows;
        f.stagesA = (int)((227 * 1024 - (long long)convf3::smem_fixed(f.chunks, f.cluster)) / ((long long)f.slab_rows * 128));
        if (f.stagesA > 4) f.stagesA = 4;
        WSI_ARG_CHECK(f.stagesA >= 2, "wsi_net_finalize: kw-split conv does not fit shared memory");
        f.relu = op.relu;
        f.has_res = op.res >= 0;
        f.bias = op.d_bias;
        f.out_c_off = op.out_c_off;
        const int K3 = 3 * f.chunks * 64;
        uint32_t es[2] = {1, 1};
        {
          uint64_t dims[2] = {(uint64_t)K3, (uint64_t)convf3::N3};
          uint64_t strides[1] = {(uint64_t)K3 * 2};
          uint32_t box[2] = {64, (uint32_t)(convf3::N3 / f.cluster)};
          int rc = wsi_make_tmap(&op.tmB, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, op.d_w, dims, strides, box, es,
                                 CU_TENSOR_MAP_SWIZZLE_128B);
          if (rc) return rc;
        }
        {
          uint64_t dims[2] = {(uint64_t)op.Cin, (uint64_t)B * ti.Hp() * ti.Wp()};
          uint64_t strides[1] = {(uint64_t)ti.C * 2};
          uint32_t box[2] = {64, (uint32_t)f.slab_box_rows};
          int rc = wsi_make_tmap(&op.tmA, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, ti.ptr + op.in_c_off, dims, strides, box, es,
                                 CU_TENSOR_MAP_SWIZZLE_128B);
          if (rc) return rc;
        }
        {
          uint64_t dims[2] = {(uint64_t)to.C, (uint64_t)B * to.Hp() * to.Wp()};
          uint64_t strides[1] = {(uint64_t)to.C * 2};
          uint32_t box[2] = {64, (uint32_t)convf3::TMO};
          int rc = wsi_make_tmap(&op.tmO, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, to.ptr, dims, strides, box, es,
                                 CU_TENSOR_MAP_SWIZZLE_128B);
          if (rc) return rc;
        }
        op.tmR = op.tmO;
        op.tmO32 = op.tmO;
        if (f.has_res) {
          const Tensor& tr = n->tensors[op.res];
          uint64_t dims[2] = {(uint64_t)op.Cout, (uint64_t)B * tr.Hp() * tr.Wp()};
          uint64_t strides[1] = {(uint64_t)tr.C * 2};
          uint32_t box[2] = {64, (uint32_t)convf3::TMO};
          int rc = wsi_make_tmap(&op.tmR, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, tr.ptr + op.res_c_off, dims, strides, box, es,
                                 CU_TENSOR_MAP_SWIZZLE_128B);
          if (rc) return rc;
        }
        int smem = (int)convf3::smem_bytes(f);
        if (smem > max_smem_flat3) max_smem_flat3 = smem;
        continue;
      }
      if (op.flat) {
        convf::Args& f = op.fa;
        memset(&f, 0, sizeof(f));
        f.Hp = ti.Hp(); f.Wp = ti.Wp(); f.H = to.H; f.W = to.W;
        f.N = op.Cout;
        f.pad = to.pad;
        f.chunks = (op.Cin + 63) / 64;
        f.ksteps = (f.chunks == 1 && op.cin_real <= 32) ? 2 : 4;
        int force_T = 0, force_nout = 0, force_res = -1;                              // tuning knobs (bench only)
        if (const char* e = getenv("WSI_FLAT_T")) force_T = atoi(e) == 1 ? 1 : 2;
        if (const char* e = getenv("WSI_FLAT_NOUT")) force_nout = atoi(e);
        if (const char* e = getenv("WSI_FLAT_RESIDENT")) force_res = atoi(e);
        // CTA pairs: each CTA keeps half of every weight k-block (N/2 rows, a whole number of 8-row swizzle groups)
        f.cluster = (n->cluster_enabled && !n->no_flat_pair && f.N % 16 == 0) ? 2 : 1;
        const int nB = f.cluster == 2 ? f.N / 2 : f.N;           // weight rows per CTA
        const bool can_reside = (size_t)9 * f.chunks * nB * 128 <= 80 * 1024;
        f.tma_epi = (op.Cout % 64 == 0 || (op.Cout % 32 == 0 && !n->legacy_epi)) ? 1 : 0;
        // shared-memory plan {T, resident weights, staging buffers per group}: prefer two M-tiles per slab (less halo
        // re-read) with resident weights; wide images (InceptionV4 stem, 150-pixel rows = 300 halo rows per slab) only
        // fit one M-tile next to resident weights, and streaming 72 KB of weights per 256 pixels costs more L2 traffic
        // than the extra halo
        const int plans[7][3] = {{2, 1, 2}, {2, 1, 1}, {1, 1, 2}, {1, 1, 1}, {2, 0, 2}, {2, 0, 1}, {1, 0, 1}};
        f.stagesA = 0;
        for (auto& pl : plans) {
          if (pl[1] && !can_reside) continue;
          if (force_T && pl[0] != force_T) continue;
          if (force_nout && pl[2] != force_nout) continue;
          if (force_res >= 0 && pl[1] != force_res) continue;
          const int T = pl[0];
          const int rows_needed = T * 128 + 2 * f.Wp + 2;
          const int loads = (rows_needed + 255) / 256;
          const int box_rows = ((rows_needed + loads - 1) / loads + 7) / 8 * 8;
          const int slab_rows = loads * box_rows;
          const int sB = pl[1] ? 9 * f.chunks : 4;
          const int nOut = f.tma_epi ? pl[2] : 0;
          const long long budget = 224 * 1024 - 2048 - (long long)sB * nB * 128 - (long long)(T * nOut) * 16384;
          int sA = (int)(budget / ((long long)slab_rows * 128));
          if (sA > 4) sA = 4;
          if (sA >= 2 && box_rows <= 256) {
            f.T = T; f.resident_b = pl[1]; f.stagesA = sA; f.stagesB = sB; f.nOut = nOut; f.nRes = 0;
            f.slab_loads = loads; f.slab_box_rows = box_rows; f.slab_rows = slab_rows;
            break;
          }
        }
        WSI_ARG_CHECK(f.stagesA >= 2 && f.slab_box_rows <= 256, "wsi_net_finalize: flat conv does not fit shared memory");
        f.relu = op.relu;
        f.has_res = op.res >= 0;
        f.res_C = op.res >= 0 ? n->tensors[op.res].C : 0;
        f.res_c_off = op.res_c_off;
        f.bias = op.d_bias;
        f.out_C = to.C; f.out_c_off = op.out_c_off;
        op.tmO = op.tmB;
        op.tmR = op.tmB;
        op.tmO32 = op.tmB;
        f.res_smem = 0;
        if (f.has_res && f.tma_epi && f.nOut == 2 && op.Cout % 64 == 0 && n->res_smem_enabled) {
          const Tensor& tr = n->tensors[op.res];
          uint64_t dims[2] = {(uint64_t)op.Cout, (uint64_t)B * tr.Hp() * tr.Wp()};
          uint64_t strides[1] = {(uint64_t)tr.C * 2};
          uint32_t box[2] = {64, 128};
          uint32_t es[2] = {1, 1};
          int rc = wsi_make_tmap(&op.tmR, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, tr.ptr + op.res_c_off, dims, strides, box, es,
                                 CU_TENSOR_MAP_SWIZZLE_128B);
          if (rc) return rc;
          f.res_smem = 1;
        }
        if (f.tma_epi) {
          uint64_t dims[2] = {(uint64_t)to.C, (uint64_t)B * to.Hp() * to.Wp()};
          uint64_t strides[1] = {(uint64_t)to.C * 2};
          uint32_t box[2] = {64, 128};
          uint32_t es[2] = {1, 1};
          int rc = WSI_OK;
          if (op.Cout >= 64)
            rc = wsi_make_tmap(&op.tmO, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, to.ptr, dims, strides, box, es,
                               CU_TENSOR_MAP_SWIZZLE_128B);
          if (rc) return rc;
          if (op.Cout & 32) {
            uint32_t box32[2] = {32, 128};
            rc = wsi_make_tmap(&op.tmO32, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, to.ptr, dims, strides, box32, es,
                               CU_TENSOR_MAP_SWIZZLE_64B);
            if (rc) return rc;
          }
        }
        int smem = (int)convf::smem_bytes(f);
        if (smem > max_smem_flat) max_smem_flat = smem;
        {
          uint64_t dims[2] = {(uint64_t)op.Ktot, (uint64_t)op.Cout};
          uint64_t strides[1] = {(uint64_t)op.Ktot * 2};
          uint32_t box[2] = {64, (uint32_t)(f.cluster == 2 ? f.N / 2 : f.N)};
          uint32_t es[2] = {1, 1};
          int rc = wsi_make_tmap(&op.tmB, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, op.d_w, dims, strides, box, es,
                                 CU_TENSOR_MAP_SWIZZLE_128B);
          if (rc) return rc;
        }
        {
          uint64_t dims[2] = {(uint64_t)op.Cin, (uint64_t)B * ti.Hp() * ti.Wp()};
          uint64_t strides[1] = {(uint64_t)ti.C * 2};
          uint32_t box[2] = {64, (uint32_t)f.slab_box_rows};
          uint32_t es[2] = {1, 1};
          int rc = wsi_make_tmap(&op.tmA, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, ti.ptr + op.in_c_off, dims, strides, box, es,
                                 CU_TENSOR_MAP_SWIZZLE_128B);
          if (rc) return rc;
        }
        continue;
      }
      convk::Args& a = op.ca;
      memset(&a, 0, sizeof(a));
      pick_box(to.W, to.H, B, &a.bw, &a.bh, &a.bn);
      a.tiles_w = wsi_cdiv(to.W, a.bw);
      a.tiles_h = wsi_cdiv(to.H, a.bh);
      a.Ho = to.H; a.Wo = to.W;
      a.stride = op.stride; a.pad_h = op.pad_h; a.pad_w = op.pad_w; a.KH = op.KH; a.KW = op.KW;
      a.cin_chunks = (op.Cin + 63) / 64;
      a.is_stem = op.is_stem;
      a.N_tile = pick_ntile(op.Cout);
      if (op.center_cols >= 0)                       // the centre-tap-only channels get N tiles of their own
        for (a.N_tile = 256; a.N_tile > 64; a.N_tile -= 64)
          if (op.center_cols % a.N_tile == 0 && (op.Cout - op.center_cols) % a.N_tile == 0) break;
      a.tiles_N = op.Cout / a.N_tile;
      a.center_nt = op.center_cols >= 0 ? op.center_cols / a.N_tile : a.tiles_N;
      a.center_kb0 = (op.pad_h * op.KW + op.pad_w) * ((op.Cin + 63) / 64);
      a.tma_epi = (a.N_tile % 64 == 0 || (a.N_tile % 32 == 0 && !n->legacy_epi)) ? 1 : 0;
      a.Cout = op.Cout;
      // CTA pairs (cta_group::2, M = 256): N must be a multiple of 32 there and each CTA's half of the weight
      // block a whole number of 8-row swizzle groups
      a.cluster = (n->cluster_enabled && !op.is_stem && a.N_tile >= 64 && a.N_tile % 32 == 0) ? 2 : 1;
      const size_t stage_bytes = convk::stage_bytes(a.N_tile, a.cluster);
      int stages = (int)((227 * 1024 - convk::smem_fixed(a.tma_epi, a.Cout)) / stage_bytes);
      if (stages > 8) stages = 8;
      if (stages < 2) stages = 2;
      a.stages = stages;
      a.out_C = to.C; a.out_c_off = op.out_c_off;
      a.res_C = op.res >= 0 ? n->tensors[op.res].C : 0;
      a.out_pad = to.pad; a.out_Hp = to.Hp(); a.out_Wp = to.Wp();
      a.res_pad = op.res >= 0 ? n->tensors[op.res].pad : 0;
      a.res_Hp = op.res >= 0 ? n->tensors[op.res].Hp() : 0;
      a.res_Wp = op.res >= 0 ? n->tensors[op.res].Wp() : 0;
      a.relu = op.relu;
      a.relu_cols = op.relu ? (op.relu_cols >= 0 ? op.relu_cols : op.Cout) : 0;
      WSI_ARG_CHECK(a.relu_cols == 0 || a.relu_cols == op.Cout || (a.tma_epi && a.relu_cols % 64 == 0),
                    "wsi_net_finalize: partial ReLU needs the TMA epilogue and a 64-channel boundary");
      a.bias = op.d_bias;
      int smem = (int)convk::smem_bytes(a.N_tile, a.stages, a.tma_epi, a.Cout, a.cluster);
      if (smem > max_smem) max_smem = smem;
      op.tmO = op.tmB;
      op.tmO32 = op.tmB;
      if (a.tma_epi) {
        // destination seen from its interior pixel (0,0) with the stored pitches: partial pixel boxes and
        // the zero border are protected by TMA's bounds clipping
        uint64_t dims[4] = {(uint64_t)to.C, (uint64_t)to.W, (uint64_t)to.H, (uint64_t)B};
        uint64_t strides[3] = {(uint64_t)to.C * 2, (uint64_t)to.Wp() * to.C * 2, (uint64_t)to.Hp() * to.Wp() * to.C * 2};
        uint32_t box[4] = {64, (uint32_t)a.bw, (uint32_t)a.bh, (uint32_t)a.bn};
        uint32_t es[4] = {1, 1, 1, 1};
        int rc = WSI_OK;
        if (a.N_tile >= 64)
          rc = wsi_make_tmap(&op.tmO, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, to.ptr + to.interior_offset(), dims, strides,
                             box, es, CU_TENSOR_MAP_SWIZZLE_128B);
        if (rc) return rc;
        op.tmR = op.tmO;
        a.res_smem = 0;
        if (op.res >= 0 && a.N_tile % 64 == 0 && n->res_smem_enabled && op.KH * op.KW * a.cin_chunks >= 16) {
          const Tensor& tr = n->tensors[op.res];
          uint64_t rdims[4] = {(uint64_t)op.Cout, (uint64_t)tr.W, (uint64_t)tr.H, (uint64_t)B};
          uint64_t rstrides[3] = {(uint64_t)tr.C * 2, (uint64_t)tr.Wp() * tr.C * 2, (uint64_t)tr.Hp() * tr.Wp() * tr.C * 2};
          rc = wsi_make_tmap(&op.tmR, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, tr.ptr + tr.interior_offset() + op.res_c_off, rdims, rstrides,
                             box, es, CU_TENSOR_MAP_SWIZZLE_128B);
          if (rc) return rc;
          a.res_smem = 1;
        }
        if (a.N_tile & 32) {
          uint32_t box32[4] = {32, (uint32_t)a.bw, (uint32_t)a.bh, (uint32_t)a.bn};
          rc = wsi_make_tmap(&op.tmO32, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, to.ptr + to.interior_offset(), dims, strides,
                             box32, es, CU_TENSOR_MAP_SWIZZLE_64B);
          if (rc) return rc;
        }
      }
      // weights: [Cout][Ktot] K-major, box 64 x N_tile, 128B swizzle
      {
        uint64_t dims[2] = {(uint64_t)op.Ktot, (uint64_t)op.Cout};
        uint64_t strides[1] = {(uint64_t)op.Ktot * 2};
        uint32_t box[2] = {(uint32_t)(op.is_stem ? 32 : 64), (uint32_t)(a.cluster == 2 ? a.N_tile / 2 : a.N_tile)};
        uint32_t es[2] = {1, 1};
        int rc = wsi_make_tmap(&op.tmB, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, op.d_w, dims, strides, box, es,
                               op.is_stem ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B);
        if (rc) return rc;
      }
      // activations: base pointer patched per forward for external inputs (map re-encoded then)
      if (!ti.external) {
        int rc = wsi_net_encode_input_map(net, (int)(&op - &n->ops[0]), ti.ptr);
        if (rc) return rc;
      }
    } else if ((op.kind == OP_MAXPOOL || op.kind == OP_AVGPOOL) && !op.skip) {
      if (!op.h_pbias.empty()) {
        WSI_CUDA_CHECK(cudaMalloc(&op.d_pbias, op.h_pbias.size() * sizeof(float)));
        WSI_CUDA_CHECK(cudaMemcpy(op.d_pbias, op.h_pbias.data(), op.h_pbias.size() * sizeof(float),
                                  cudaMemcpyHostToDevice));
      }
      // 3x3 pools over internal tensors: tile plan + input tensor map of the TMA-ring kernel
      const Tensor& ti = n->tensors[op.in];
      const Tensor& to = n->tensors[op.out];
      const int cv = (op.pC ? op.pC : ti.C) / 8;
      op.pt_ct = 0;
      if (op.pk == 3 && (op.pstride == 1 || op.pstride == 2) && ti.C % 8 == 0) {
        // stride 1: a thread owns two adjacent output columns (k_pool3s1_tma), so a slot is a column PAIR
        const int per = op.pstride == 1 ? 2 : 1;
        const int units_w = (to.W + per - 1) / per;
        int ct = 0, wt = 0;
        for (int c = 32; c >= 4; --c)                   // whole image width per CTA when <= 288 slots; inner TMA
          if (cv % c == 0 && c * units_w <= 288 && c * units_w > ct * ((wt + per - 1) / per)) { ct = c; wt = to.W; }   // rows >= 64 B
        if (!ct) {
          ct = cv % 8 == 0 ? 8 : (cv % 4 == 0 ? 4 : (cv % 2 == 0 ? 2 : 1));
          const int cap = 288 / ct * per;               // output columns per CTA
          const int ntiles = (to.W + cap - 1) / cap;
          wt = (to.W + ntiles - 1) / ntiles;
          wt = (wt + per - 1) / per * per;
        }
        const int cols = per == 2 ? (wt + 1) / 2 * 2 + 2 : (wt - 1) * op.pstride + 3;
        const int row_stride = (cols * ct * 16 + 127) / 128 * 128;
        int R = (40 * 1024) / row_stride;
        if (R > 12) R = 12;
        if (cols <= 256 && R >= 3) {
          op.pt_ct = ct; op.pt_wt = wt; op.pt_cols = cols; op.pt_R = R; op.pt_per = per;
          if (!ti.external) {                           // an external input is mapped at every forward
            int rc = encode_pool_map(n, op, ti.ptr);
            if (rc) return rc;
          }
        }
      }
    } else if (op.kind == OP_HEAD) {
      WSI_CUDA_CHECK(cudaMalloc(&op.d_fcw, op.h_fcw.size() * sizeof(float)));
      WSI_CUDA_CHECK(cudaMemcpy(op.d_fcw, op.h_fcw.data(), op.h_fcw.size() * sizeof(float), cudaMemcpyHostToDevice));
      WSI_CUDA_CHECK(cudaMalloc(&op.d_fcb, op.h_fcb.size() * sizeof(float)));
      WSI_CUDA_CHECK(cudaMemcpy(op.d_fcb, op.h_fcb.data(), op.h_fcb.size() * sizeof(float), cudaMemcpyHostToDevice));
    }
  }
  if (max_smem > 0)
    WSI_CUDA_CHECK(cudaFuncSetAttribute(convk::k_conv_igemm<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem));
  if (max_smem > 0)
    WSI_CUDA_CHECK(cudaFuncSetAttribute(convk::k_conv_igemm<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem));
  if (max_smem_stem > 0)
    WSI_CUDA_CHECK(cudaFuncSetAttribute(convs::k_conv_stem<false, 7, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem_stem));
  if (max_smem_stem > 0)
    WSI_CUDA_CHECK(cudaFuncSetAttribute(convs::k_conv_stem<false, 3, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem_stem));
  if (max_smem_stem_pool > 0)
    WSI_CUDA_CHECK(cudaFuncSetAttribute(convs::k_conv_stem<true, 7, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem_stem_pool));
  if (max_smem_flat > 0) {
    WSI_CUDA_CHECK(cudaFuncSetAttribute(convf::k_conv_flat<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem_flat));
    WSI_CUDA_CHECK(cudaFuncSetAttribute(convf::k_conv_flat<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem_flat));
  }
  if (max_smem_flat3 > 0) {
    WSI_CUDA_CHECK(cudaFuncSetAttribute(convf3::k_conv_flat3<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem_flat3));
    WSI_CUDA_CHECK(cudaFuncSetAttribute(convf3::k_conv_flat3<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem_flat3));
  }
  WSI_CUDA_CHECK(cudaFuncSetAttribute(k_pool3s1_tma<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
  WSI_CUDA_CHECK(cudaFuncSetAttribute(k_pool3s1_tma<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
  WSI_CUDA_CHECK(cudaFuncSetAttribute(k_pool3_tma<true, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
  WSI_CUDA_CHECK(cudaFuncSetAttribute(k_pool3_tma<true, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
  WSI_CUDA_CHECK(cudaFuncSetAttribute(k_pool3_tma<false, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
  WSI_CUDA_CHECK(cudaFuncSetAttribute(k_pool3_tma<false, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
  n->launches_per_forward = 0;
  for (auto& op : n->ops) n->launches_per_forward += op.skip ? 0 : 1;
  n->finalized = true;
  return WSI_OK;
}

// (re)encode the A-operand tensor map of conv op `op_index` for input base pointer `base`
int wsi_net_encode_input_map(void* net, int op_index, const void* base) {
  Net* n = (Net*)net;
  WSI_ARG_CHECK(n && op_index >= 0 && op_index < (int)n->ops.size(), "wsi_net_encode_input_map: bad op");
  Op& op = n->ops[op_index];
  WSI_ARG_CHECK(op.kind == OP_CONV && base, "wsi_net_encode_input_map: not a conv");
  const Tensor& ti = n->tensors[op.in];
  const convk::Args& a = op.ca;
  const uint64_t B = (uint64_t)n->max_batch;
  if (op.is_stem) {
    // overlapping-window view of the padded NHWC4 image: element (k, wo, row, n) lives at
    // base + n*H*rowb + row*rowb + wo*16 + k*2  (k < 32: 8 pixels x 4 channels from pixel 2*wo)
    const uint64_t rowb = (uint64_t)ti.W * 8;   // 4 ch x bf16
    uint64_t dims[4] = {32, (uint64_t)a.Wo, (uint64_t)ti.H, B};
    uint64_t strides[3] = {16, rowb, (uint64_t)ti.H * rowb};
    uint32_t box[4] = {32, (uint32_t)a.bw, (uint32_t)(a.bh * 2), (uint32_t)a.bn};
    uint32_t es[4] = {1, 1, 2, 1};
    return wsi_make_tmap(&op.tmA, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, base, dims, strides, box, es,
                         CU_TENSOR_MAP_SWIZZLE_64B);
  }
  // a padded tensor is addressed from its interior pixel (0,0) with the padded pitches; TMA's
  // out-of-bounds zero fill then stands in for the border exactly as for a dense tensor
  base = (const __nv_bfloat16*)base + ti.interior_offset() + op.in_c_off;
  uint64_t dims[4] = {(uint64_t)op.Cin, (uint64_t)ti.W, (uint64_t)ti.H, B};
  uint64_t strides[3] = {(uint64_t)ti.C * 2, (uint64_t)ti.Wp() * ti.C * 2, (uint64_t)ti.Hp() * ti.Wp() * ti.C * 2};
  uint32_t box[4] = {64, (uint32_t)(a.bw * a.stride), (uint32_t)(a.bh * a.stride), (uint32_t)a.bn};
  uint32_t es[4] = {1, (uint32_t)a.stride, (uint32_t)a.stride, 1};
  return wsi_make_tmap(&op.tmA, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, base, dims, strides, box, es,
                       CU_TENSOR_MAP_SWIZZLE_128B);
}

int wsi_net_forward(void* net, const void* d_input, int B, float* d_probs, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  Net* n = (Net*)net;
  WSI_ARG_CHECK(n && n->finalized, "wsi_net_forward: net not finalized");
  WSI_ARG_CHECK(B >= 1 && B <= n->max_batch, "wsi_net_forward: batch %d out of range (max %d)", B, n->max_batch);
  cudaEvent_t* pev = nullptr;
  if (n->prof_slots > 0)
    pev = n->prof_ev.data() + (size_t)(n->prof_forwards % n->prof_slots) * (n->ops.size() + 1);
  for (size_t i = 0; i < n->ops.size(); ++i) {
    Op& op = n->ops[i];
    if (pev) cudaEventRecord(pev[i], stream);
    if (op.skip) continue;                         // pooling executed inside the stem kernel
    const Tensor& ti = n->tensors[op.in];
    const __nv_bfloat16* in_ptr = ti.external ? (const __nv_bfloat16*)d_input : ti.ptr;
    WSI_ARG_CHECK(in_ptr, "wsi_net_forward: missing input pointer");
    if (op.kind == OP_CONV) {
      if (ti.external && op.is_stem != 2) {
        int rc = wsi_net_encode_input_map(net, (int)i, in_ptr);
        if (rc) return rc;
      }
      const Tensor& to = n->tensors[op.out];
      if (op.is_stem == 2) {
        convs::Args sa = op.sa;
        sa.in = (const uint8_t*)in_ptr;
        sa.total_q = (long long)B * sa.P * sa.P;
        sa.num_tiles = (int)((sa.total_q + convs::T * 128 - 1) / (convs::T * 128));
        if (op.fused_pool >= 0) {
          sa.B = B;
          int grid = B < n->num_sms ? B : n->num_sms;
          cudaLaunchConfig_t cfg;
          memset(&cfg, 0, sizeof(cfg));
          cfg.gridDim = dim3((unsigned)grid);
          cfg.blockDim = dim3(convs::THREADS_POOL);
          cfg.dynamicSmemBytes = convs::smem_bytes(sa.slab_pairs, sa.stages, true);
          cfg.stream = stream;
          cudaLaunchAttribute attr[1];
          attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
          attr[0].val.programmaticStreamSerializationAllowed = 1;
          cfg.attrs = attr;
          cfg.numAttrs = n->pdl_enabled ? 1 : 0;
          WSI_CUDA_CHECK(cudaLaunchKernelEx(&cfg, convs::k_conv_stem<true, 7, 2>, op.tmB, sa));
          continue;
        }
        int grid = sa.num_tiles < n->num_sms ? sa.num_tiles : n->num_sms;
        if (op.KH == 7)
          convs::k_conv_stem<false, 7, 2><<<grid, convs::THREADS, convs::smem_bytes(sa.slab_pairs, sa.stages), stream>>>(op.tmO, sa);
        else
          convs::k_conv_stem<false, 3, 1><<<grid, convs::THREADS,
                                            convs::smem_bytes(sa.slab_pairs, sa.stages, false, 3 * op.Cout * 32), stream>>>(op.tmO, sa);
        continue;
      }
      if (op.flat3) {
        convf3::Args f = op.f3;
        f.B = B;
        f.total_rows = (long long)B * f.Hp * f.Wp;
        f.num_tiles = (int)((f.total_rows + convf3::TMO - 1) / convf3::TMO);
        if (f.cluster == 2) {
          const long long units = (f.num_tiles + 1) / 2;
          long long pairs = units < n->num_sms / 2 ? units : n->num_sms / 2;
          cudaLaunchConfig_t cfg;
          memset(&cfg, 0, sizeof(cfg));
          cfg.blockDim = dim3(convf3::THREADS);
          cfg.dynamicSmemBytes = convf3::smem_bytes(f);
          cfg.stream = stream;
          cudaLaunchAttribute attr[2];
          attr[0].id = cudaLaunchAttributeClusterDimension;
          attr[0].val.clusterDim.x = 2;
          attr[0].val.clusterDim.y = 1;
          attr[0].val.clusterDim.z = 1;
          attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;   // prologue overlaps the previous conv's tail
          attr[1].val.programmaticStreamSerializationAllowed = 1;
          cfg.attrs = attr;
          cfg.numAttrs = n->pdl_enabled ? 2 : 1;
          if (op.max_pairs == 0) {
            int nclusters = 0;
            cfg.gridDim = dim3((unsigned)(n->num_sms / 2 * 2));
            if (cudaOccupancyMaxActiveClusters(&nclusters, convf3::k_conv_flat3<2>, &cfg) != cudaSuccess || nclusters <= 0) {
              cudaGetLastError();
              nclusters = n->num_sms / 2;
            }
            op.max_pairs = nclusters;
          }
          if (pairs > op.max_pairs) pairs = op.max_pairs;
          cfg.gridDim = dim3((unsigned)(2 * pairs));
          WSI_CUDA_CHECK(cudaLaunchKernelEx(&cfg, convf3::k_conv_flat3<2>, op.tmA, op.tmB, op.tmO, op.tmR, f));
          continue;
        }
        int grid = f.num_tiles < n->num_sms ? f.num_tiles : n->num_sms;
        convf3::k_conv_flat3<1><<<grid, convf3::THREADS, convf3::smem_bytes(f), stream>>>(op.tmA, op.tmB, op.tmO, op.tmR, f);
        continue;
      }
      if (op.flat) {
        convf::Args f = op.fa;
        f.B = B;
        f.total_rows = (long long)B * f.Hp * f.Wp;
        f.num_tiles = (int)((f.total_rows + f.T * 128 - 1) / (f.T * 128));
        f.out = to.ptr;
        f.residual = op.res >= 0 ? n->tensors[op.res].ptr + op.res_c_off : nullptr;
        if (f.cluster == 2) {
          const long long units = (f.num_tiles + 1) / 2;
          long long pairs = units < n->num_sms / 2 ? units : n->num_sms / 2;
          cudaLaunchConfig_t cfg;
          memset(&cfg, 0, sizeof(cfg));
          cfg.blockDim = dim3(convf::THREADS);
          cfg.dynamicSmemBytes = convf::smem_bytes(f);
          cfg.stream = stream;
          cudaLaunchAttribute attr[2];
          attr[0].id = cudaLaunchAttributeClusterDimension;
          attr[0].val.clusterDim.x = 2;
          attr[0].val.clusterDim.y = 1;
          attr[0].val.clusterDim.z = 1;
          attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;   // prologue overlaps the previous conv's tail
          attr[1].val.programmaticStreamSerializationAllowed = 1;
          cfg.attrs = attr;
          cfg.numAttrs = n->pdl_enabled ? 2 : 1;
          if (op.max_pairs == 0) {
            // GPCs with an odd SM count cannot host a pair on every SM: size the persistent grid to what is co-resident
            int nclusters = 0;
            cfg.gridDim = dim3((unsigned)(n->num_sms / 2 * 2));
            if (cudaOccupancyMaxActiveClusters(&nclusters, convf::k_conv_flat<2>, &cfg) != cudaSuccess || nclusters <= 0) {
              cudaGetLastError();
              nclusters = n->num_sms / 2;
            }
            op.max_pairs = nclusters;
          }
          if (pairs > op.max_pairs) pairs = op.max_pairs;
          cfg.gridDim = dim3((unsigned)(2 * pairs));
          WSI_CUDA_CHECK(cudaLaunchKernelEx(&cfg, convf::k_conv_flat<2>, op.tmA, op.tmB, op.tmO, op.tmO32, op.tmR, f));
          continue;
        }
        int grid = f.num_tiles < n->num_sms ? f.num_tiles : n->num_sms;
        convf::k_conv_flat<1><<<grid, convf::THREADS, convf::smem_bytes(f), stream>>>(op.tmA, op.tmB, op.tmO, op.tmO32, op.tmR, f);
        continue;
      }
      convk::Args a = op.ca;
      a.B = B;
      a.tiles_n = wsi_cdiv(B, a.bn);
      a.out = to.ptr;
      a.residual = op.res >= 0 ? n->tensors[op.res].ptr + op.res_c_off : nullptr;
      const long long total = (long long)a.tiles_w * a.tiles_h * a.tiles_n * a.tiles_N;
      int grid = (int)(total < n->num_sms ? total : n->num_sms);
      size_t smem = convk::smem_bytes(a.N_tile, a.stages, a.tma_epi, a.Cout, a.cluster);
      if (a.cluster == 2) {
        // CTA pairs: an even grid of 2-CTA clusters, one pair per two M tiles
        const long long m_tiles = (long long)a.tiles_w * a.tiles_h * a.tiles_n;
        const long long units = (m_tiles + 1) / 2 * a.tiles_N;
        long long pairs = units < n->num_sms / 2 ? units : n->num_sms / 2;
        cudaLaunchConfig_t cfg;
        memset(&cfg, 0, sizeof(cfg));
        cfg.gridDim = dim3((unsigned)(2 * pairs));
        cfg.blockDim = dim3(convk::THREADS);
        cfg.dynamicSmemBytes = smem;
        cfg.stream = stream;
        cudaLaunchAttribute attr[2];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = 2;
        attr[0].val.clusterDim.y = 1;
        attr[0].val.clusterDim.z = 1;
        attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;   // prologue overlaps the previous conv's tail
        attr[1].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr;
        cfg.numAttrs = n->pdl_enabled ? 2 : 1;
        if (op.max_pairs == 0) {
          // GPCs with an odd SM count cannot host a pair on every SM: size the persistent grid to what is co-resident
          int nclusters = 0;
          cudaLaunchConfig_t q = cfg;
          q.gridDim = dim3((unsigned)(n->num_sms / 2 * 2));
          if (cudaOccupancyMaxActiveClusters(&nclusters, convk::k_conv_igemm<2>, &q) != cudaSuccess || nclusters <= 0) {
            cudaGetLastError();
            nclusters = n->num_sms / 2;
          }
          op.max_pairs = nclusters;
        }
        if (pairs > op.max_pairs) pairs = op.max_pairs;
        cfg.gridDim = dim3((unsigned)(2 * pairs));
        WSI_CUDA_CHECK(cudaLaunchKernelEx(&cfg, convk::k_conv_igemm<2>, op.tmA, op.tmB, op.tmO, op.tmO32, op.tmR, a));
      } else {
        convk::k_conv_igemm<1><<<grid, convk::THREADS, smem, stream>>>(op.tmA, op.tmB, op.tmO, op.tmO32, op.tmR, a);
      }
    } else if (op.kind == OP_MAXPOOL || op.kind == OP_AVGPOOL) {
      const Tensor& to = n->tensors[op.out];
      const bool sliced = op.in_c_off != 0 || (op.pC != 0 && op.pC != ti.C);
      const int cv = (op.pC ? op.pC : ti.C) / 8;
      const bool is_max = op.kind == OP_MAXPOOL;
      WSI_ARG_CHECK(!sliced || op.pt_ct > 0, "wsi_net_forward: a pool over a channel slice needs the TMA-ring plan");
      if (op.pk == 3 && (op.pstride == 1 || op.pstride == 2)) {
        if ((n->pool_ring || sliced) && op.pt_ct > 0) {
          if (ti.external) {
            int rc = encode_pool_map(n, op, in_ptr);
            if (rc) return rc;
          }
          const int row_bytes = op.pt_cols * op.pt_ct * 16, row_stride = (row_bytes + 127) / 128 * 128;
          const size_t smem = 256 + (size_t)op.pt_R * row_stride;
          dim3 grid(B, (to.W + op.pt_wt - 1) / op.pt_wt, cv / op.pt_ct);
          const int threads = ((op.pt_wt + op.pt_per - 1) / op.pt_per * op.pt_ct + 31) / 32 * 32;
          if (op.pt_per == 2) {
            if (is_max)
              k_pool3s1_tma<true><<<grid, threads, smem, stream>>>(op.tmA, ti.H, ti.W, op.ppad, op.count_include_pad, to.H,
                                                                   to.W, to.ptr, to.C, op.out_c_off, ti.Hp(), ti.pad,
                                                                   to.Hp(), to.Wp(), to.pad, op.pt_ct, op.pt_wt, row_bytes,
                                                                   row_stride, op.pt_R, nullptr, 0);
            else
              k_pool3s1_tma<false><<<grid, threads, smem, stream>>>(op.tmA, ti.H, ti.W, op.ppad, op.count_include_pad, to.H,
                                                                    to.W, to.ptr, to.C, op.out_c_off, ti.Hp(), ti.pad,
                                                                    to.Hp(), to.Wp(), to.pad, op.pt_ct, op.pt_wt, row_bytes,
                                                                    row_stride, op.pt_R, op.d_pbias, op.prelu);
            continue;
          }
#define WSI_POOL_TMA(MAX, S)                                                                                             \
  k_pool3_tma<MAX, S><<<grid, threads, smem, stream>>>(op.tmA, ti.H, ti.W, op.ppad, op.count_include_pad, to.H, to.W,    \
                                                       to.ptr, to.C, op.out_c_off, ti.Hp(), ti.pad, to.Hp(), to.Wp(),    \
                                                       to.pad, op.pt_ct, op.pt_wt, row_bytes, row_stride, op.pt_R)
          if (is_max) {
            if (op.pstride == 1) WSI_POOL_TMA(true, 1); else WSI_POOL_TMA(true, 2);
          } else {
            if (op.pstride == 1) WSI_POOL_TMA(false, 1); else WSI_POOL_TMA(false, 2);
          }
#undef WSI_POOL_TMA
          continue;
        }
        const int nbands = to.H > 24 ? (to.H + 15) / 16 : 1;     // <= 16 output rows per CTA, bands of equal height
        const int band = (to.H + nbands - 1) / nbands;
        dim3 block(256), grid(B * nbands, (to.W * cv + 255) / 256);
#define WSI_POOL_SEP(MAX, S)                                                                                              \
  k_pool_sep<MAX, 3, S><<<grid, block, 0, stream>>>(in_ptr, ti.H, ti.W, ti.C, op.ppad, op.count_include_pad, to.H, to.W, \
                                                    to.ptr, to.C, op.out_c_off, ti.Hp(), ti.Wp(), ti.pad, to.Hp(),       \
                                                    to.Wp(), to.pad, band, nbands, op.d_pbias, op.prelu)
        if (is_max) {
          if (op.pstride == 1) WSI_POOL_SEP(true, 1); else WSI_POOL_SEP(true, 2);
        } else {
          if (op.pstride == 1) WSI_POOL_SEP(false, 1); else WSI_POOL_SEP(false, 2);
        }
#undef WSI_POOL_SEP
      } else {
        const int bx = cv < 256 ? cv : 256;
        int by = 256 / bx;
        if (by > to.W) by = to.W;
        dim3 block(bx, by);
        if (is_max)
          k_pool<true><<<B * to.H, block, 0, stream>>>(in_ptr, ti.H, ti.W, ti.C, op.pk, op.pstride, op.ppad, 0, to.H, to.W,
                                                       to.ptr, to.C, op.out_c_off, ti.Hp(), ti.Wp(), ti.pad, to.Hp(), to.Wp(), to.pad);
        else
          k_pool<false><<<B * to.H, block, 0, stream>>>(in_ptr, ti.H, ti.W, ti.C, op.pk, op.pstride, op.ppad,
                                                        op.count_include_pad, to.H, to.W, to.ptr, to.C, op.out_c_off, ti.Hp(),
                                                        ti.Wp(), ti.pad, to.Hp(), to.Wp(), to.pad);
      }
    } else if (op.kind == OP_HEAD) {
      WSI_ARG_CHECK(d_probs, "wsi_net_forward: null probs");
      k_head<<<B, HEAD_THREADS, ti.C * sizeof(float), stream>>>(in_ptr, ti.H * ti.W, ti.C, op.d_fcw, op.d_fcb, op.K,
                                                                d_probs);
    }
  }
  if (pev) {
    cudaEventRecord(pev[n->ops.size()], stream);
    n->prof_forwards++;
  }
  wsi_count_launch(n->launches_per_forward);
  WSI_CUDA_CHECK(cudaGetLastError());
  return WSI_OK;
}

int wsi_net_set_profiling(void* net, int slots) {
  Net* n = (Net*)net;
  WSI_ARG_CHECK(n && n->finalized && slots >= 0, "wsi_net_set_profiling: bad arguments");
  for (auto e : n->prof_ev) cudaEventDestroy(e);
  n->prof_ev.clear();
  n->prof_slots = slots;
  n->prof_forwards = 0;
  n->prof_ev.resize((size_t)slots * (n->ops.size() + 1));
  for (auto& e : n->prof_ev) WSI_CUDA_CHECK(cudaEventCreate(&e));
  return WSI_OK;
}

// Sum of per-op durations (ms) over the forwards recorded since wsi_net_set_profiling (at most `slots`
// most recent); synchronises the device.  h_op_kind: 0 conv, 1 maxpool, 2 avgpool, 3 head.
int wsi_net_profile_read(void* net, double* h_op_ms, double* h_op_flops, int* h_op_kind, int max_ops, int* h_forwards) {
  Net* n = (Net*)net;
  WSI_ARG_CHECK(n && h_op_ms && h_forwards, "wsi_net_profile_read: bad arguments");
  WSI_CUDA_CHECK(cudaDeviceSynchronize());
  const int nops = (int)n->ops.size();
  WSI_ARG_CHECK(max_ops >= nops, "wsi_net_profile_read: need room for %d ops", nops);
  const long long cnt = n->prof_forwards < n->prof_slots ? n->prof_forwards : n->prof_slots;
  for (int i = 0; i < nops; ++i) {
    h_op_ms[i] = 0;
    if (h_op_flops) h_op_flops[i] = n->ops[i].flops_per_image;
    if (h_op_kind) h_op_kind[i] = (int)n->ops[i].kind;
  }
  for (long long s = 0; s < cnt; ++s) {
    cudaEvent_t* ev = n->prof_ev.data() + (size_t)s * (nops + 1);
    for (int i = 0; i < nops; ++i) {
      float ms = 0;
      WSI_CUDA_CHECK(cudaEventElapsedTime(&ms, ev[i], ev[i + 1]));
      h_op_ms[i] += ms;
    }
  }
  *h_forwards = (int)cnt;
  return WSI_OK;
}

int wsi_net_tensor_ptr(void* net, int id, void** ptr, int* hwc3) {
  Net* n = (Net*)net;
  WSI_ARG_CHECK(n && id >= 0 && id < (int)n->tensors.size() && ptr, "wsi_net_tensor_ptr: bad id");
  *ptr = n->tensors[id].ptr;
  if (hwc3) {
    hwc3[0] = n->tensors[id].H; hwc3[1] = n->tensors[id].W; hwc3[2] = n->tensors[id].C;
  }
  return WSI_OK;
}

// copy `batch` images between a dense bf16 [batch][H][W][C] device buffer and tensor `id` (whatever its padding)
static int copy_tensor(void* net, int id, int batch, void* d_dense, int to_tensor, void* stream_) {
  Net* n = (Net*)net;
  WSI_ARG_CHECK(n && n->finalized && id >= 0 && id < (int)n->tensors.size() && d_dense, "wsi_net_read/write_tensor: bad arguments");
  const Tensor& t = n->tensors[id];
  WSI_ARG_CHECK(t.ptr && batch >= 1 && batch <= n->max_batch && t.C % 8 == 0, "wsi_net_read/write_tensor: bad tensor");
  size_t total = (size_t)batch * t.H * t.W * (t.C / 8);
  int blocks = (int)((total + 255) / 256);
  if (to_tensor)
    k_copy_interior<<<blocks, 256, 0, (cudaStream_t)stream_>>>((const uint4*)d_dense, (uint4*)t.ptr, batch, t.H, t.W, t.C / 8, t.Hp(), t.Wp(), t.pad, 1);
  else
    k_copy_interior<<<blocks, 256, 0, (cudaStream_t)stream_>>>((const uint4*)t.ptr, (uint4*)d_dense, batch, t.H, t.W, t.C / 8, t.Hp(), t.Wp(), t.pad, 0);
  WSI_CUDA_CHECK(cudaGetLastError());
  return WSI_OK;
}

int wsi_net_read_tensor(void* net, int id, int batch, void* d_dense_out, void* stream) {
  return copy_tensor(net, id, batch, d_dense_out, 0, stream);
}

int wsi_net_write_tensor(void* net, int id, int batch, const void* d_dense_in, void* stream) {
  return copy_tensor(net, id, batch, (void*)d_dense_in, 1, stream);
}

// 1 if conv op `op_index` runs on the flat-slab kernel (convf::k_conv_flat)
int wsi_net_op_is_flat(void* net, int op_index) {
  Net* n = (Net*)net;
  if (!n || op_index < 0 || op_index >= (int)n->ops.size()) return 0;
  return n->ops[op_index].flat ? 1 : 0;
}

int wsi_net_input_geometry(void* net, int* hwc3) {
  Net* n = (Net*)net;
  WSI_ARG_CHECK(n && hwc3, "wsi_net_input_geometry: null pointer");
  for (auto& t : n->tensors)
    if (t.external) {
      hwc3[0] = t.H; hwc3[1] = t.W; hwc3[2] = t.C;
      return WSI_OK;
    }
  wsi_set_error("wsi_net_input_geometry: the network has no external input tensor");
  return WSI_ERR_ARG;
}

// info[0] = layout kind of the external input (0: bf16 [B][H][W][4] with the patch at the stem's padding offset;
// 1: two row-parity planes of 16-byte pixel pairs, bf16 [2][max_batch][P][P][8], see conv_stem.cuh),
// info[1] = P, info[2..3] = plane stride in PIXELS (low, high 32 bits)
int wsi_net_input_layout(void* net, int* info4) {
  Net* n = (Net*)net;
  WSI_ARG_CHECK(n && info4, "wsi_net_input_layout: null pointer");
  info4[0] = info4[1] = info4[2] = info4[3] = 0;
  for (auto& op : n->ops)
    if (op.kind == OP_CONV && op.is_stem == 2 && n->tensors[op.in].external) {
      const int P = n->tensors[op.in].H / 2;
      const long long stride_px = (long long)n->max_batch * P * P * 2;
      info4[0] = 1; info4[1] = P;
      info4[2] = (int)(stride_px & 0xffffffffll); info4[3] = (int)(stride_px >> 32);
    }
  return WSI_OK;
}

// bytes the caller must allocate for the external input of a max_batch forward (includes the slack the
// stem's bulk copies may read past the last image)
size_t wsi_net_input_bytes(void* net) {
  Net* n = (Net*)net;
  if (!n) return 0;
  for (auto& t : n->tensors)
    if (t.external) return (size_t)n->max_batch * t.H * t.W * t.C * sizeof(__nv_bfloat16) + 32 * 1024;
  return 0;
}

int wsi_net_num_classes(void* net) {
  Net* n = (Net*)net;
  if (n)
    for (auto& op : n->ops)
      if (op.kind == OP_HEAD) return op.K;
  return 0;
}

int wsi_net_num_ops(void* net) {
  Net* n = (Net*)net;
  return n ? (int)n->ops.size() : 0;
}

int wsi_net_num_launches(void* net) {
  Net* n = (Net*)net;
  return n ? n->launches_per_forward : 0;
}

double wsi_net_flops_per_image(void* net) {
  Net* n = (Net*)net;
  double f = 0;
  if (n)
    for (auto& op : n->ops) f += op.flops_per_image;
  return f;
}

// Algorithmic HBM bytes of the convolution launches per image: every conv reads its (slice of the) input once,
// its residual once, its weights once per forward (amortised over the batch: not counted) and writes its output
// once; a stem with the max-pooling fused writes the pooled tensor instead of the conv output.
double wsi_net_conv_bytes_per_image(void* net) {
  Net* n = (Net*)net;
  double b = 0;
  if (!n) return 0;
  for (auto& op : n->ops) {
    if (op.kind != OP_CONV) continue;
    const Tensor& ti = n->tensors[op.in];
    const Tensor* to = &n->tensors[op.out];
    if (op.fused_pool >= 0) to = &n->tensors[n->ops[op.fused_pool].out];
    const double cin = op.is_stem ? (double)ti.C : (double)op.Cin;
    b += 2.0 * ti.H * ti.W * cin + 2.0 * to->H * to->W * op.Cout;
    if (op.res >= 0) b += 2.0 * to->H * to->W * op.Cout;
  }
  return b;
}

void wsi_net_destroy(void* net) {
  Net* n = (Net*)net;
  if (!n) return;
  for (auto e : n->prof_ev) cudaEventDestroy(e);
  for (auto& t : n->tensors)
    if (t.ptr) cudaFree(t.ptr);
  for (auto& op : n->ops) {
    if (op.d_w) cudaFree(op.d_w);
    if (op.d_bias) cudaFree(op.d_bias);
    if (op.d_fcw) cudaFree(op.d_fcw);
    if (op.d_pbias) cudaFree(op.d_pbias);
    if (op.d_fcb) cudaFree(op.d_fcb);
  }
  delete n;
}

}  // extern "C"
