// K4b: the 7x7 / stride 2 / pad 3 stem convolution (3 -> Cout=64 channels) as a tcgen05 implicit GEMM
// that reads its overlapping input windows straight from UNEXPANDED image rows in shared memory.
//
// The per-tap kernel (conv.cuh) fed the stem through an "overlapping window" tensor map: every
// output pixel's 8-pixel window was materialised by TMA, 64 bytes at a time, three times over
// (ncu: 10 GB L2->SM per launch for a 0.43 GB input, tensor pipe 12 % active, 1.5 ms).  Here:
//
//  * the preprocess kernel writes the padded (S+6)^2 x 4-channel image as two ROW-PARITY planes of
//    16-byte pixel pairs, bf16 [2][maxB][P][P][8] with P = (S+6)/2:  element [s][n][r][x][e] is
//    padded row 2r+s, column 2x + e/4, channel e%4 (channel 3 = 0);
//  * in flat pair index q = (n*P + ho)*P + wo the window of output (ho, wo) for tap row kh = 2a+s is
//    the 64 bytes starting at pair q + a*P of plane s -- a pure shift, and consecutive outputs are
//    consecutive pairs, i.e. operand rows 16 bytes apart that overlap by 48 bytes;
//  * that is exactly a K-major NO-SWIZZLE UMMA operand whose core matrices overlap: row pitch 16 B
//    inside a core matrix, K-direction core stride (LBO) 16 B, 8-row-group stride (SBO) 128 B
//    (tools/umma_noswz_probe.cu, profiles/r01_umma_noswz_probe.txt: bit-exact on B200).
//
// One 1-D bulk copy per parity plane brings (T*128 + 3P + 4) pairs (9.7 KB) for T*128 outputs; the 14
// k-steps (7 tap rows x 2 pair-pairs, K = 16 each) are MMAs whose A start address moves by a*P + 2*jj
// pairs.  Weights (28 KB, canonical no-swizzle blocks) stay resident.  Outputs are produced for all P
// columns of a row (3 of 115 are junk that nobody reads) into a [maxB*P*P][Cout] tensor.
//
// Warps: 0 producer, 1 MMA issuer, 2-5 epilogue group 0 (M-tile 0), 6-9 epilogue group 1 (M-tile 1);
// each group stages its 128x64 tile in swizzled smem and writes it with one TMA store.
#pragma once
#include "conv.cuh"

namespace convs {

constexpr int THREADS = 320;
constexpr int THREADS_POOL = 576;   // POOL: 16 epilogue warps -- (M-tile, 32-column half) x 4 TMEM lane quarters
constexpr int T = 2;             // M-tiles per super-tile
// Instantiations: KH = 7, JJ = 2 (7x7/2 pad 3: 7 tap rows x 2 pair-pairs = 14 K steps of 16) and KH = 3, JJ = 1
// (3x3/2 pad 0, InceptionV4 conv2d_1a: the 3 taps of a row are pixels 2wo..2wo+2 = pairs wo, wo+1 -> one K step
// per tap row).  Cout (UMMA N) is 64 or 32 at run time.
constexpr int N_MAX = 64;

struct Args {
  int P;                         // pairs per padded row == padded rows per parity plane
  int N;                         // Cout: 64 | 32
  long long total_q;             // B * P * P flat outputs
  long long plane_pairs;         // maxB * P * P: distance between the two parity planes (pairs)
  int num_tiles;
  int slab_pairs;                // pairs per parity slab (multiple of 8)
  int stages;
  int relu;
  // fused 3x3 / stride 2 / pad 1 max pooling (POOL instantiation): tiles are output-row PAIRS of one image
  int B, row_pairs, pool_w;      // images, Ho/2 tiles per image, pooled columns
  __nv_bfloat16* pool_out;       // pooled tensor [B][pool_Hp][pool_Wp][pool_C], interior at (pool_pad, pool_pad)
  int pool_Hp, pool_Wp, pool_pad, pool_C, pool_c_off;
  const uint8_t* in;             // parity-plane input (bf16 pairs)
  const uint8_t* w;              // weights in canonical no-swizzle blocks [14][8][2][8][8] bf16
  const float* bias;
};

__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
          ptx::smem_u32(dst)),
      "l"(reinterpret_cast<uint64_t>(src)), "r"(bytes), "r"(ptx::smem_u32(bar))
      : "memory");
}

__device__ __forceinline__ uint32_t convk_hmax2(uint32_t a, uint32_t b) {
  __nv_bfloat162 r = __hmax2(*reinterpret_cast<__nv_bfloat162*>(&a), *reinterpret_cast<__nv_bfloat162*>(&b));
  return *reinterpret_cast<uint32_t*>(&r);
}

__device__ __forceinline__ void umma_lohi2(uint32_t tmem_d, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo, uint32_t b_hi,
                                           uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\t"
      "mov.b64 da, {%1, %2};\n\t"
      "mov.b64 db, {%3, %4};\n\t"
      "setp.ne.b32 p, %6, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %5, p;\n\t}"
      ::"r"(tmem_d), "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate)
      : "memory");
}

// 16 fp32 accumulator columns -> + bias -> ReLU -> 8 packed bf16 pairs
__device__ __forceinline__ void bias_relu_pack16(const uint32_t (&d)[16], const float* s_bias, uint32_t (&p)[8]) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float4 b = *reinterpret_cast<const float4*>(s_bias + 4 * i);
    p[2 * i] = convk::pack_bf16(fmaxf(__uint_as_float(d[4 * i]) + b.x, 0.f), fmaxf(__uint_as_float(d[4 * i + 1]) + b.y, 0.f));
    p[2 * i + 1] = convk::pack_bf16(fmaxf(__uint_as_float(d[4 * i + 2]) + b.z, 0.f), fmaxf(__uint_as_float(d[4 * i + 3]) + b.w, 0.f));
  }
}

// POOL = false: one super-tile = T*128 consecutive flat outputs, each written to the conv output tensor.
// POOL = true : ReLU + the 3x3/2 max pooling that follows the stem in every ResNet are fused into the epilogue
// (the unfused pair wrote and re-read 1.7 GB per 1024 patches: 0.48 + 0.66 ms of an 7.8 ms step).  A CTA walks
// whole images; tile r of an image is output rows 2r (M-tile 0) and 2r+1 (M-tile 1), BOTH starting at column 0: the
// second MMA's A window starts one image row (P pairs) below the first instead of 128 positions further, so TMEM lane
// l of the two accumulators holds the two vertically adjacent outputs of column l (needs P <= 128).
// 16 epilogue warps (4 TMEM lane quarters x 4 sixteen-channel quarters): a thread loads its column's 16 channels of
// both rows, adds bias, ReLUs, packs to bf16 and takes the VERTICAL maximum of rows 2r-1 (carried in registers from the
// previous tile), 2r, 2r+1 without touching shared memory; only that one row is staged (16 KB, double-buffered), and
// after a named barrier 448 threads each reduce one pooled (column, channel octet) from three staged columns and store
// it straight into the (padded) pooled tensor.  Post-ReLU values are >= 0, so a missing neighbour is a 0; the maximum
// commutes with the (monotone) bf16 rounding, so the result equals pooling the rounded conv output.
// (Tried and dropped: 8 warps that only drain + 6 warps that only pool, handing the buffers over through named
// barriers -- 0.66 ms against 0.44 ms: what the drain needs is MANY warps in flight, each is a dependent chain.)
// History (ncu): 8 epilogue warps staging both rows and reading 9 positions per output: 529 us, tensor pipe 40 % active,
// latency-bound; 16 warps + row 2r+1's horizontal maximum carried: 452 us, 48 %, 79 % of the shared-memory data pipe.
template <bool POOL, int KH, int JJ>
__global__ void __launch_bounds__(POOL ? THREADS_POOL : THREADS, 1) k_conv_stem(const __grid_constant__ CUtensorMap tmapO, const Args a) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem = smem_raw + (smem_base - ptx::smem_u32(smem_raw));
  constexpr int KSTEPS = KH * JJ;
  const int N = a.N;
  const int W_BYTES = KSTEPS * N * 32;
  const int slab_bytes = a.slab_pairs * 16;           // one parity
  uint8_t* sOut = smem;                               // 4 x 16 KB: [group][buffer], 1024-aligned (POOL: 2 x 16 KB)
  uint8_t* sW = sOut + (POOL ? 2 : 4) * 16384;
  uint8_t* sA = sW + W_BYTES;                          // stages x 2 parities x slab_bytes
  float* sBias = reinterpret_cast<float*>(sA + (size_t)a.stages * 2 * slab_bytes);
  uint64_t* bars = reinterpret_cast<uint64_t*>(sBias + N_MAX);
  uint64_t* full = bars;
  uint64_t* empty = full + a.stages;
  uint64_t* tfull = empty + a.stages;
  uint64_t* tempty = tfull + 2;
  uint64_t* wfull = tempty + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(wfull + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0 && lane == 0) {
    if (!POOL) ptx::prefetch_tmap(&tmapO);
    for (int s = 0; s < a.stages; ++s) {
      ptx::mbar_init(&full[s], 1);
      ptx::mbar_init(&empty[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      ptx::mbar_init(&tfull[s], 1);
      ptx::mbar_init(&tempty[s], POOL ? 16 : 8);
    }
    ptx::mbar_init(wfull, 1);
    ptx::fence_barrier_init();
  }
  if (warp == 1) {
    ptx::tmem_alloc(tmem_slot, 256);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // barriers, tensor memory and descriptors are set up: only now wait for the producing kernel of the stream (its
  // tail overlapped this prologue), and let the consuming kernel do the same with ours
  ptx::pdl_wait();
  ptx::pdl_launch_dependents();

  if (warp == 0) {
    // ===================== producer: weights once, then one bulk copy per parity per super-tile =====================
    if (ptx::elect_one()) {
      ptx::mbar_arrive_expect_tx(wfull, (uint32_t)W_BYTES);
      for (int i = 0; i < 2; ++i) bulk_g2s(sW + i * (W_BYTES / 2), a.w + i * (W_BYTES / 2), W_BYTES / 2, wfull);
    }
    __syncwarp();
    int stage = 0;
    uint32_t phase = 0;
    const int my_tiles = POOL ? ((a.B - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x) * a.row_pairs
                              : (a.num_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
    for (int it = 0; it < my_tiles; ++it) {
      ptx::mbar_wait(&empty[stage], phase ^ 1);
      if (ptx::elect_one()) {
        long long q0;
        if (POOL) {
          const int n = (int)blockIdx.x + (it / a.row_pairs) * (int)gridDim.x, r = it % a.row_pairs;
          q0 = ((long long)n * a.P + 2 * r) * a.P;
        } else {
          q0 = (long long)((int)blockIdx.x + it * (int)gridDim.x) * (T * 128);
        }
        ptx::mbar_arrive_expect_tx(&full[stage], (uint32_t)(2 * slab_bytes));
        uint8_t* dst = sA + (size_t)stage * 2 * slab_bytes;
        bulk_g2s(dst, a.in + q0 * 16, (uint32_t)slab_bytes, &full[stage]);
        bulk_g2s(dst + slab_bytes, a.in + (a.plane_pairs + q0) * 16, (uint32_t)slab_bytes, &full[stage]);
      }
      __syncwarp();
      if (++stage == a.stages) { stage = 0; phase ^= 1; }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    const uint32_t idesc = ptx::make_idesc_bf16(128, N);
    // no-swizzle K-major descriptors: LBO (K-direction core stride) in bits 16-29, SBO (8-row group) in 32-45
    const uint32_t a_hi = (uint32_t)(128 >> 4) | (1u << 14);
    const uint32_t a_lbo = (uint32_t)(16 >> 4) << 16;
    const uint32_t b_hi = (uint32_t)(256 >> 4) | (1u << 14);
    const uint32_t b_lbo = (uint32_t)(128 >> 4) << 16;
    const uint32_t w_lo = ((ptx::smem_u32(sW) & 0x3FFFFu) >> 4) | b_lbo;
    const uint32_t sA_lo = ((ptx::smem_u32(sA) & 0x3FFFFu) >> 4) | a_lbo;
    const uint32_t stage_step = (uint32_t)(2 * slab_bytes) >> 4, par_step = (uint32_t)slab_bytes >> 4;
    const uint32_t P = (uint32_t)a.P;
    ptx::mbar_wait(wfull, 0);
    int stage = 0, acc = 0;
    uint32_t phase = 0, pacc = 0;
    const int my_tiles = POOL ? ((a.B - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x) * a.row_pairs
                              : (a.num_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
    for (int it = 0; it < my_tiles; ++it) {
      ptx::mbar_wait(&tempty[acc], pacc ^ 1);
      ptx::mbar_wait(&full[stage], phase);
      ptx::tc_fence_after();
      if (ptx::elect_one()) {
        const uint32_t d0 = tmem_base + (uint32_t)(acc * T * N);
        const uint32_t s_lo = sA_lo + (uint32_t)stage * stage_step;
#pragma unroll
        for (int ks = 0; ks < KSTEPS; ++ks) {
          const int kh = ks / JJ, jj = ks % JJ;
          // pair offset inside the parity slab: a*P rows down, 2*jj pairs right (units of 16 B == descriptor units)
          const uint32_t a_lo = s_lo + (uint32_t)(kh & 1) * par_step + (uint32_t)(kh >> 1) * P + (uint32_t)(2 * jj);
          const uint32_t b_lo = w_lo + (uint32_t)ks * (uint32_t)(N * 32 >> 4);
          umma_lohi2(d0, a_lo, a_hi, b_lo, b_hi, idesc, ks == 0 ? 0u : 1u);
          umma_lohi2(d0 + (uint32_t)N, a_lo + (POOL ? P : 128u), a_hi, b_lo, b_hi, idesc, ks == 0 ? 0u : 1u);
        }
        ptx::umma_commit(&empty[stage]);
        ptx::umma_commit(&tfull[acc]);
      }
      __syncwarp();
      if (++stage == a.stages) { stage = 0; phase ^= 1; }
      if (++acc == 2) { acc = 0; pacc ^= 1; }
    }
  } else {
    // ===================== epilogue: group g = M-tile g of every super-tile =====================
    const int g = (warp - 2) >> 2;                    // !POOL: M-tile 0 | 1.  POOL: 16-channel quarter 0..3
    const int q = warp & 3;                           // TMEM lane quarter of this warp
    const int row = q * 32 + lane;                    // !POOL: flat position in the M-tile.  POOL: output column
    const int gt = ((warp - 2) & 3) * 32 + lane;      // 0..127 inside the group
    const bool leader = gt == 0;
    if ((int)threadIdx.x - 64 < N) sBias[(int)threadIdx.x - 64] = a.bias[(int)threadIdx.x - 64];
    ptx::named_bar_sync(3, POOL ? 512 : 256);         // bias staged (all epilogue warps)
    if (POOL) {
      const int et = (int)threadIdx.x - 64;             // 0..511
      const int my_imgs = (a.B - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
      const int items = a.pool_w * 8;                   // (pooled column, 16-byte channel octet)
      const uint32_t sbuf0 = ptx::smem_u32(sOut);
      auto ldcol = [&](uint32_t buf, uint32_t col, uint32_t oct) {
        return ptx::ld_shared_v4(buf + col * 128u + ((oct ^ (col & 7u)) << 4));
      };
      auto vmax = [](uint4& m, const uint4& x) {
        __nv_bfloat162* mm = reinterpret_cast<__nv_bfloat162*>(&m);
        const __nv_bfloat162* xx = reinterpret_cast<const __nv_bfloat162*>(&x);
#pragma unroll
        for (int i = 0; i < 4; ++i) mm[i] = __hmax2(mm[i], xx[i]);
      };
      int acc = 0, bsel = 0;
      uint32_t pacc = 0;
      uint32_t carry[8];                                // row 2r-1 of this column, 16 channels (bf16 pairs)
      for (int im = 0; im < my_imgs; ++im) {
        const int n = (int)blockIdx.x + im * (int)gridDim.x;
        __nv_bfloat16* oimg = a.pool_out + (((size_t)n * a.pool_Hp + a.pool_pad) * a.pool_Wp + a.pool_pad) * a.pool_C + a.pool_c_off;
        for (int r = 0; r < a.row_pairs; ++r) {
          const uint32_t buf = sbuf0 + (uint32_t)bsel * 16384u;
          ptx::mbar_wait(&tfull[acc], pacc);
          ptx::tc_fence_after();
          const uint32_t t0 = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * T * N + g * 16);
          uint32_t d0[16], d1[16], p0[8], p1[8];
          ptx::tmem_ld_32x32b_x16(t0, d0);                          // row 2r
          ptx::tmem_ld_32x32b_x16(t0 + (uint32_t)N, d1);            // row 2r + 1
          ptx::tmem_ld_wait();
          ptx::tc_fence_before();
          __syncwarp();
          if (lane == 0) ptx::mbar_arrive(&tempty[acc]);
          bias_relu_pack16(d0, sBias + g * 16, p0);
          bias_relu_pack16(d1, sBias + g * 16, p1);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            uint32_t m = convk_hmax2(p0[i], p1[i]);
            if (r > 0) m = convk_hmax2(m, carry[i]);
            carry[i] = p1[i];
            p0[i] = m;
          }
          const uint32_t rb = buf + (uint32_t)row * 128u, sw = (uint32_t)row & 7u;
          ptx::st_shared_v4(rb + ((((uint32_t)(2 * g)) ^ sw) << 4), p0[0], p0[1], p0[2], p0[3]);
          ptx::st_shared_v4(rb + ((((uint32_t)(2 * g + 1)) ^ sw) << 4), p0[4], p0[5], p0[6], p0[7]);
          ptx::named_bar_sync(1, 512);                  // vertical maxima of all 128 columns x 64 channels staged
          if (et < items) {
            const uint32_t pc = (uint32_t)et >> 3, oct = (uint32_t)et & 7u;
            uint4 m = ldcol(buf, 2u * pc, oct);
            vmax(m, ldcol(buf, 2u * pc + 1u, oct));
            if (pc > 0) vmax(m, ldcol(buf, 2u * pc - 1u, oct));
            *reinterpret_cast<uint4*>(oimg + ((size_t)r * a.pool_Wp + pc) * a.pool_C + oct * 8) = m;
          }
          bsel ^= 1;
          if (++acc == 2) { acc = 0; pacc ^= 1; }
        }
      }
    } else {
    uint8_t* myOut = sOut + (size_t)g * 2 * 16384;
    int acc = 0, k = 0;
    uint32_t pacc = 0;
    for (int tile = blockIdx.x; tile < a.num_tiles; tile += gridDim.x, ++k) {
      const int ob = k & 1;
      if (leader) ptx::bulk_wait_read<1>();
      ptx::named_bar_sync(1 + g, 128);                // staging buffer ob of this group is free again
      ptx::mbar_wait(&tfull[acc], pacc);
      ptx::tc_fence_after();
      const uint32_t t_addr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)((acc * T + g) * N);
      if (N == 64) {
        convk::epi_chunk64(t_addr, row, true, ptx::smem_u32(myOut + (size_t)ob * 16384), 0u, sBias, a.relu);
      } else {
        const uint4 none[8] = {};
        convk::epi_chunk32_regs(t_addr, row, ptx::smem_u32(myOut + (size_t)ob * 16384), false, none, sBias, a.relu);
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&tempty[acc]);
      ptx::fence_proxy_async();
      ptx::named_bar_sync(1 + g, 128);                // tile staged
      if (leader) {
        ptx::tma_store_2d(&tmapO, myOut + (size_t)ob * 16384, 0, (int)((long long)tile * (T * 128) + g * 128));
        ptx::bulk_commit();
      }
      if (++acc == 2) { acc = 0; pacc ^= 1; }
    }
    if (leader) ptx::bulk_wait<0>();
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    __syncwarp();
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, 256);
  }
}

static inline size_t smem_bytes(int slab_pairs, int stages, bool pool = false, int w_bytes = 14 * 64 * 32) {
  return 1024 + (pool ? 2 : 4) * 16384 + w_bytes + (size_t)stages * 2 * slab_pairs * 16 + 256 + (size_t)(2 * stages + 5) * 8 + 16;
}

}  // namespace convs
