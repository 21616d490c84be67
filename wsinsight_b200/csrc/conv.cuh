// K4a/K4b: convolution as a tcgen05/TMEM implicit GEMM fed by TMA (sm_100a).
//
// Replaces the cuDNN convolutions inside the reference's TorchScript forward
// (wsinsight/modellib/run_inference.py:542-545).  Activations are NHWC bf16,
// weights bf16 K-major [Cout][KH*KW*Cin_pad], accumulation fp32 in TMEM.
//
// GEMM view: M = output pixels (tile = 128 pixels = bw x bh x bn box of the output
// tensor), N = Cout (UMMA N <= 256), K = taps x channels, BLOCK_K = 64 channels of
// one tap.  The A tile of one (tap, channel-chunk) is ONE 4-D TMA box of the input
// tensor: start = (c0, wo0*s + kw - pad, ho0*s + kh - pad, n0), traversal strides
// (1,s,s,1); out-of-bounds coordinates are zero-filled by TMA == conv zero padding.
// No im2col buffer ever exists in HBM.
//
// Warp roles (192 threads, 1 CTA/SM, persistent over tiles):
//   warp 0  : TMA producer        (smem ring, full/empty mbarriers)
//   warp 1  : TMEM alloc + MMA issuer (tcgen05.mma, one elected lane)
//   warps 2-5: epilogue            (tcgen05.ld -> +bias (+residual) -> ReLU -> bf16 -> NHWC store)
// Two TMEM accumulator stages overlap the epilogue of tile i with the MMAs of tile i+1.
#pragma once
#include "common.cuh"

namespace convk {

constexpr int BLOCK_M = 128;
constexpr int BLOCK_K = 64;
constexpr int THREADS = 320;   // warp 0 producer, warp 1 MMA, warps 2-5 / 6-9 two epilogue groups
constexpr int A_BYTES = BLOCK_M * BLOCK_K * 2;   // 16 KB

struct Args {
  // tiling of the output tensor
  int tiles_w, tiles_h, tiles_n, tiles_N;   // total tiles = tiles_w*tiles_h*tiles_n*tiles_N
  int bw, bh, bn;                            // output pixels per M tile along w, h, n (bw*bh*bn == 128)
  int Ho, Wo, B;
  int stride, pad_h, pad_w, KH, KW, cin_chunks;
  int is_stem;                               // 7x7/2 stem: overlapping-window map, 7 K-chunks of 32 (SWIZZLE_64B)
  int N_tile;                                // UMMA N (multiple of 32, <= 256)
  int stages;
  int out_C, out_c_off;                      // channels of the destination tensor / concat offset
  int res_C;                                 // channels of the residual tensor (0 = none)
  int res_smem;                              // residual boxes are TMA-loaded into the staging buffers (tmapR)
  int out_pad, res_pad;                      // one-pixel zero border of the destination / residual tensor (0|1)
  int out_Hp, out_Wp, res_Hp, res_Wp;        // stored rows / columns of the destination / residual tensor
  int tma_epi;                               // N_tile % 64 == 0: smem-staged epilogue + TMA stores, two warp groups
  int Cout;                                  // tiles_N * N_tile (bias entries staged in smem)
  int cluster;                               // 1, or 2: CTA PAIRS (cta_group::2).  One tcgen05.mma of the leader CTA
                                             // spans both SMs: M = 256 (two M tiles of the same N tile), each CTA
                                             // stages its own 128 activation rows and HALF of the weight k-block, so a
                                             // stage is 16 KB + N_tile*64 B and the ring is 1.5x deeper (the single-CTA
                                             // ring starved: ncu showed the MMA warp waiting on TMA data 26 % of the time)
  int center_nt, center_kb0;                 // N tiles >= center_nt run only the cin_chunks k-blocks from center_kb0
                                             // (channels whose weights are zero outside the centre tap); tiles_N = none
  int relu;
  int relu_cols;                             // TMA epilogue: ReLU on output channels < relu_cols (merged sibling convs
                                             // may end with a slice that has none); Cout or 0 for ordinary convs
  const __nv_bfloat16* residual;
  const float* bias;                         // [tiles_N * N_tile]
  __nv_bfloat16* out;
};

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float bf16_lo(uint32_t v) { return __uint_as_float(v << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t v) { return __uint_as_float(v & 0xffff0000u); }

// Epilogue of one 64-column chunk of a 128-row accumulator tile, staged for a TMA store:
// TMEM -> registers -> (+bias, +residual from its smem staging, ReLU) -> bf16 -> shared memory in the
// SWIZZLE_128B layout of a [128 rows][64 ch] box (16-byte chunk c of row r lives at r*128 + ((c ^ (r&7))<<4)),
// so that a quarter warp covers all 32 banks and ONE bulk tensor store writes the tile with full
// 128-byte lines (the per-thread 16-byte stores at a row pitch it replaces were the bottleneck of
// every layer with N <= 128).  `keep` false writes zeros (border rows of the padded flat layout).
__device__ __forceinline__ void epi_chunk64(uint32_t t_addr, int row, bool keep, uint32_t s_out, uint32_t s_res,
                                            const float* s_bias, int relu) {
  uint32_t r[64];
  ptx::tmem_ld_32x32b_x32(t_addr, r);
  ptx::tmem_ld_32x32b_x32(t_addr + 32u, r + 32);
  ptx::tmem_ld_wait();
  const uint32_t row_base = (uint32_t)row * 128u;
  const uint32_t sw = (uint32_t)row & 7u;
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    const float4 b0 = *reinterpret_cast<const float4*>(s_bias + 8 * c);
    const float4 b1 = *reinterpret_cast<const float4*>(s_bias + 8 * c + 4);
    float v0 = __uint_as_float(r[8 * c + 0]) + b0.x, v1 = __uint_as_float(r[8 * c + 1]) + b0.y;
    float v2 = __uint_as_float(r[8 * c + 2]) + b0.z, v3 = __uint_as_float(r[8 * c + 3]) + b0.w;
    float v4 = __uint_as_float(r[8 * c + 4]) + b1.x, v5 = __uint_as_float(r[8 * c + 5]) + b1.y;
    float v6 = __uint_as_float(r[8 * c + 6]) + b1.z, v7 = __uint_as_float(r[8 * c + 7]) + b1.w;
    const uint32_t off = row_base + (((uint32_t)c ^ sw) << 4);
    if (s_res) {
      const uint4 rv = ptx::ld_shared_v4(s_res + off);
      v0 += bf16_lo(rv.x); v1 += bf16_hi(rv.x); v2 += bf16_lo(rv.y); v3 += bf16_hi(rv.y);
      v4 += bf16_lo(rv.z); v5 += bf16_hi(rv.z); v6 += bf16_lo(rv.w); v7 += bf16_hi(rv.w);
    }
    if (relu) {
      v0 = fmaxf(v0, 0.f); v1 = fmaxf(v1, 0.f); v2 = fmaxf(v2, 0.f); v3 = fmaxf(v3, 0.f);
      v4 = fmaxf(v4, 0.f); v5 = fmaxf(v5, 0.f); v6 = fmaxf(v6, 0.f); v7 = fmaxf(v7, 0.f);
    }
    uint32_t o0 = pack_bf16(v0, v1), o1 = pack_bf16(v2, v3), o2 = pack_bf16(v4, v5), o3 = pack_bf16(v6, v7);
    if (!keep) o0 = o1 = o2 = o3 = 0u;
    ptx::st_shared_v4(s_out + off, o0, o1, o2, o3);
  }
}

// Same, with the residual already in registers (8 x 16 B = this thread's 64 channels): used where shared
// memory has no room for a residual staging ring; the caller issues the global loads one chunk ahead.
__device__ __forceinline__ void epi_chunk64_regs(uint32_t t_addr, int row, uint32_t s_out, bool has_res,
                                                 const uint4 (&rv)[8], const float* s_bias, int relu,
                                                 bool keep = true) {
  uint32_t r[64];
  ptx::tmem_ld_32x32b_x32(t_addr, r);
  ptx::tmem_ld_32x32b_x32(t_addr + 32u, r + 32);
  ptx::tmem_ld_wait();
  const uint32_t row_base = (uint32_t)row * 128u;
  const uint32_t sw = (uint32_t)row & 7u;
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    const float4 b0 = *reinterpret_cast<const float4*>(s_bias + 8 * c);
    const float4 b1 = *reinterpret_cast<const float4*>(s_bias + 8 * c + 4);
    float v0 = __uint_as_float(r[8 * c + 0]) + b0.x, v1 = __uint_as_float(r[8 * c + 1]) + b0.y;
    float v2 = __uint_as_float(r[8 * c + 2]) + b0.z, v3 = __uint_as_float(r[8 * c + 3]) + b0.w;
    float v4 = __uint_as_float(r[8 * c + 4]) + b1.x, v5 = __uint_as_float(r[8 * c + 5]) + b1.y;
    float v6 = __uint_as_float(r[8 * c + 6]) + b1.z, v7 = __uint_as_float(r[8 * c + 7]) + b1.w;
    if (has_res) {
      v0 += bf16_lo(rv[c].x); v1 += bf16_hi(rv[c].x); v2 += bf16_lo(rv[c].y); v3 += bf16_hi(rv[c].y);
      v4 += bf16_lo(rv[c].z); v5 += bf16_hi(rv[c].z); v6 += bf16_lo(rv[c].w); v7 += bf16_hi(rv[c].w);
    }
    if (relu) {
      v0 = fmaxf(v0, 0.f); v1 = fmaxf(v1, 0.f); v2 = fmaxf(v2, 0.f); v3 = fmaxf(v3, 0.f);
      v4 = fmaxf(v4, 0.f); v5 = fmaxf(v5, 0.f); v6 = fmaxf(v6, 0.f); v7 = fmaxf(v7, 0.f);
    }
    uint32_t o0 = pack_bf16(v0, v1), o1 = pack_bf16(v2, v3), o2 = pack_bf16(v4, v5), o3 = pack_bf16(v6, v7);
    if (!keep) o0 = o1 = o2 = o3 = 0u;   // border rows of the padded flat layout stay zero
    ptx::st_shared_v4(s_out + row_base + (((uint32_t)c ^ sw) << 4), o0, o1, o2, o3);
  }
}

// Same, with the residual tile already IN the staging buffer (TMA-loaded in the swizzled layout the output uses):
// every thread reads its own 8 x 16 B, adds, and writes the result back to the same addresses.  A thread's row is a
// TMEM lane, so residual loads from global memory touch 32 different 128-byte lines per warp instruction; through
// TMA + conflict-free LDS the same bytes cost 1/8 of the L1/shared wavefronts (layer1 of ResNet34: 0.30 -> see
// DESIGN.md 4.1).
__device__ __forceinline__ void epi_chunk64_smemres(uint32_t t_addr, int row, uint32_t s_buf, const float* s_bias,
                                                    int relu, bool keep = true) {
  uint32_t r[64];
  ptx::tmem_ld_32x32b_x32(t_addr, r);
  ptx::tmem_ld_32x32b_x32(t_addr + 32u, r + 32);
  const uint32_t row_base = s_buf + (uint32_t)row * 128u;
  const uint32_t sw = (uint32_t)row & 7u;
  ptx::tmem_ld_wait();
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    const uint4 rv = ptx::ld_shared_v4(row_base + (((uint32_t)c ^ sw) << 4));
    const float4 b0 = *reinterpret_cast<const float4*>(s_bias + 8 * c);
    const float4 b1 = *reinterpret_cast<const float4*>(s_bias + 8 * c + 4);
    float v0 = __uint_as_float(r[8 * c + 0]) + b0.x, v1 = __uint_as_float(r[8 * c + 1]) + b0.y;
    float v2 = __uint_as_float(r[8 * c + 2]) + b0.z, v3 = __uint_as_float(r[8 * c + 3]) + b0.w;
    float v4 = __uint_as_float(r[8 * c + 4]) + b1.x, v5 = __uint_as_float(r[8 * c + 5]) + b1.y;
    float v6 = __uint_as_float(r[8 * c + 6]) + b1.z, v7 = __uint_as_float(r[8 * c + 7]) + b1.w;
    v0 += bf16_lo(rv.x); v1 += bf16_hi(rv.x); v2 += bf16_lo(rv.y); v3 += bf16_hi(rv.y);
    v4 += bf16_lo(rv.z); v5 += bf16_hi(rv.z); v6 += bf16_lo(rv.w); v7 += bf16_hi(rv.w);
    if (relu) {
      v0 = fmaxf(v0, 0.f); v1 = fmaxf(v1, 0.f); v2 = fmaxf(v2, 0.f); v3 = fmaxf(v3, 0.f);
      v4 = fmaxf(v4, 0.f); v5 = fmaxf(v5, 0.f); v6 = fmaxf(v6, 0.f); v7 = fmaxf(v7, 0.f);
    }
    uint32_t o0 = pack_bf16(v0, v1), o1 = pack_bf16(v2, v3), o2 = pack_bf16(v4, v5), o3 = pack_bf16(v6, v7);
    if (!keep) o0 = o1 = o2 = o3 = 0u;   // border rows of the padded flat layout stay zero
    ptx::st_shared_v4(row_base + (((uint32_t)c ^ sw) << 4), o0, o1, o2, o3);
  }
}

// 32-column tail chunk (N_tile = 64a + 32: Cout 32, 96, 160, 224, 448 of the Inception family): the same
// epilogue staged as [128 rows][32 ch] = 64-byte rows in the SWIZZLE_64B layout of a {32 ch} TMA store box
// (16-byte chunk c of row r at r*64 + ((c ^ ((r >> 1) & 3)) << 4); conflict-free for a quarter warp).
__device__ __forceinline__ void epi_chunk32_regs(uint32_t t_addr, int row, uint32_t s_out, bool has_res,
                                                 const uint4 (&rv)[8], const float* s_bias, int relu,
                                                 bool keep = true) {
  uint32_t r[32];
  ptx::tmem_ld_32x32b_x32(t_addr, r);
  ptx::tmem_ld_wait();
  const uint32_t row_base = (uint32_t)row * 64u;
  const uint32_t sw = ((uint32_t)row >> 1) & 3u;
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    const float4 b0 = *reinterpret_cast<const float4*>(s_bias + 8 * c);
    const float4 b1 = *reinterpret_cast<const float4*>(s_bias + 8 * c + 4);
    float v0 = __uint_as_float(r[8 * c + 0]) + b0.x, v1 = __uint_as_float(r[8 * c + 1]) + b0.y;
    float v2 = __uint_as_float(r[8 * c + 2]) + b0.z, v3 = __uint_as_float(r[8 * c + 3]) + b0.w;
    float v4 = __uint_as_float(r[8 * c + 4]) + b1.x, v5 = __uint_as_float(r[8 * c + 5]) + b1.y;
    float v6 = __uint_as_float(r[8 * c + 6]) + b1.z, v7 = __uint_as_float(r[8 * c + 7]) + b1.w;
    if (has_res) {
      v0 += bf16_lo(rv[c].x); v1 += bf16_hi(rv[c].x); v2 += bf16_lo(rv[c].y); v3 += bf16_hi(rv[c].y);
      v4 += bf16_lo(rv[c].z); v5 += bf16_hi(rv[c].z); v6 += bf16_lo(rv[c].w); v7 += bf16_hi(rv[c].w);
    }
    if (relu) {
      v0 = fmaxf(v0, 0.f); v1 = fmaxf(v1, 0.f); v2 = fmaxf(v2, 0.f); v3 = fmaxf(v3, 0.f);
      v4 = fmaxf(v4, 0.f); v5 = fmaxf(v5, 0.f); v6 = fmaxf(v6, 0.f); v7 = fmaxf(v7, 0.f);
    }
    uint32_t o0 = pack_bf16(v0, v1), o1 = pack_bf16(v2, v3), o2 = pack_bf16(v4, v5), o3 = pack_bf16(v6, v7);
    if (!keep) o0 = o1 = o2 = o3 = 0u;
    ptx::st_shared_v4(s_out + row_base + (((uint32_t)c ^ sw) << 4), o0, o1, o2, o3);
  }
}

// CL = 1: one CTA per tile.  CL = 2: CTA pairs (cta_group::2); a separate instantiation because a kernel that
// contains cta_group::2 instructions can only be launched as 2-CTA clusters.
template <int CL>
__global__ void __launch_bounds__(THREADS, 1)
    k_conv_igemm(const __grid_constant__ CUtensorMap tmapA, const __grid_constant__ CUtensorMap tmapB,
                 const __grid_constant__ CUtensorMap tmapO, const __grid_constant__ CUtensorMap tmapO32,
                 const __grid_constant__ CUtensorMap tmapR, const Args a) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // carve: [stages x A 16KB][stages x B N_tile*128][barriers]
  if (ptx::smem_u32(smem_raw) & 1023u) __trap();
  uint8_t* smem = smem_raw;
  const int B_BYTES = (CL == 2 ? a.N_tile / 2 : a.N_tile) * BLOCK_K * 2;   // per-CTA stage stride (the stem uses half)
  const int kbytes = a.is_stem ? 64 : 128;              // bytes of K per smem row
  uint8_t* sA = smem;
  uint8_t* sB = smem + (size_t)a.stages * A_BYTES;
  uint8_t* sOut = sB + (size_t)a.stages * B_BYTES;                       // tma_epi: 2 x 16 KB (one per group)
  float* sBias = reinterpret_cast<float*>(sOut + (a.tma_epi ? 2 * 16384 : 0));
  uint64_t* bars = reinterpret_cast<uint64_t*>(sBias + (a.tma_epi ? ((a.Cout + 31) & ~31) : 0));
  uint64_t* full = bars;
  uint64_t* empty = bars + a.stages;
  uint64_t* tfull = bars + 2 * a.stages;
  uint64_t* tempty = tfull + 2;
  uint64_t* resFull = tempty + 2;                  // residual tile landed in the staging buffer of group g
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(resFull + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m_tiles = a.tiles_w * a.tiles_h * a.tiles_n;
  // work units: one (M tile, N tile) per CTA and iteration.  In pair mode unit u of the PAIR is
  // (M-tile pair u / tiles_N, N tile u % tiles_N) and CTA `crank` takes M tile 2*pair + crank, clamped to the
  // last tile (an odd tile count makes the last pair compute the same tile twice: identical stores).
  const uint32_t crank = CL == 2 ? ptx::cluster_ctarank() : 0u;
  const int m_units = CL == 2 ? (m_tiles + 1) / 2 : m_tiles;
  const int total_units = m_units * a.tiles_N;
  const int first_unit = CL == 2 ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int unit_step = CL == 2 ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  // iteration i of this CTA (pair) -> (N tile, M tile); false when the CTA has run out of work.  Default order: unit
  // u = first + i * step, N tile fastest (neighbouring CTAs share an A tile through L2 at the same time).  With
  // centre-tap-only N tiles (center_nt < tiles_N: a few k-blocks instead of all) a CTA instead takes ALL N tiles of its
  // M tile back to back, so every CTA gets the same mix of long and short units.
  const bool by_mtile = a.center_nt < a.tiles_N;
  auto unit_of_iter = [&](int i, int& nt, int& mt) -> bool {
    if (by_mtile) {
      mt = first_unit + (i / a.tiles_N) * unit_step;
      if (mt >= m_units) return false;
      nt = i % a.tiles_N;
    } else {
      const int u = first_unit + i * unit_step;
      if (u >= total_units) return false;
      nt = u % a.tiles_N;
      mt = u / a.tiles_N;
    }
    if (CL == 2) mt = min(2 * mt + (int)crank, m_tiles - 1);
    return true;
  };
  const int num_kb = a.is_stem ? 7 : a.KH * a.KW * a.cin_chunks;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tmapA);
    ptx::prefetch_tmap(&tmapB);
    for (int s = 0; s < a.stages; ++s) {
      ptx::mbar_init(&full[s], 1);
      ptx::mbar_init(&empty[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      ptx::mbar_init(&tfull[s], 1);
      ptx::mbar_init(&tempty[s], (uint32_t)(((a.tma_epi && a.N_tile > 64) ? 8 : 4) * CL));   // > 64: two chunks or more
      ptx::mbar_init(&resFull[s], 1);
    }
    ptx::fence_barrier_init();
  }
  if (CL == 2) {
    __syncthreads();
    ptx::cluster_sync();                      // both CTAs' barriers exist before the pair-wide TMEM allocation
  }
  if (warp == 1) {
    if (CL == 2) {
      ptx::tmem_alloc_2sm(tmem_slot, 512);
      ptx::tmem_relinquish_2sm();
    } else {
      ptx::tmem_alloc(tmem_slot, 512);
      ptx::tmem_relinquish();
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (CL == 2) ptx::cluster_sync();   // the peer's barriers are initialised before anything remote happens
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // barriers, tensor memory and descriptors are set up: only now wait for the producing kernel of the stream (its
  // tail overlapped this prologue), and let the consuming kernel do the same with ours
  ptx::pdl_wait();
  ptx::pdl_launch_dependents();

  if (warp == 0) {
    // ===================== TMA producer (whole warp loops, one elected lane issues) =====================
    int stage = 0;
    uint32_t phase = 0;
    for (int it = 0;; ++it) {
      int nt, mt;
      if (!unit_of_iter(it, nt, mt)) break;
      const int tw = mt % a.tiles_w;
      mt /= a.tiles_w;
      const int th = mt % a.tiles_h;
      const int tn = mt / a.tiles_h;
      const int wo0 = tw * a.bw, ho0 = th * a.bh, n0 = tn * a.bn;
      const int kb0 = nt >= a.center_nt ? a.center_kb0 : 0, kb1 = nt >= a.center_nt ? kb0 + a.cin_chunks : num_kb;
      for (int kb = kb0; kb < kb1; ++kb) {
        ptx::mbar_wait(&empty[stage], phase ^ 1);
        if (ptx::elect_one()) {
          // pair mode: only the leader arms its barrier, for the bytes of BOTH CTAs (2 x (16 KB + half B))
          if (CL != 2 || crank == 0)
            ptx::mbar_arrive_expect_tx(&full[stage], (uint32_t)((CL == 2 ? 2 * BLOCK_M : BLOCK_M) + a.N_tile) * kbytes);
          void* dA = sA + (size_t)stage * A_BYTES;
          void* dB = sB + (size_t)stage * B_BYTES;
          if (CL == 2) {
            // own 128 activation rows + own half of the weight k-block; completion lands on the LEADER's barrier
            const int cc = kb % a.cin_chunks;
            const int tap = kb / a.cin_chunks;
            const int kw = tap % a.KW, kh = tap / a.KW;
            const uint32_t lbar = ptx::leader_bar_addr(&full[stage]);
            ptx::tma_load_4d_2sm(dA, &tmapA, lbar, cc * BLOCK_K, wo0 * a.stride + kw - a.pad_w,
                                 ho0 * a.stride + kh - a.pad_h, n0);
            ptx::tma_load_2d_2sm(dB, &tmapB, lbar, kb * BLOCK_K, nt * a.N_tile + (int)crank * (a.N_tile >> 1));
          } else
          if (a.is_stem) {
            // rows 2*ho + kh of the padded image, 8 pixels x 4 channels starting at pixel 2*wo
            ptx::tma_load_4d(dA, &tmapA, &full[stage], 0, wo0, 2 * ho0 + kb, n0);
            ptx::tma_load_2d(dB, &tmapB, &full[stage], kb * 32, nt * a.N_tile);
          } else {
            const int cc = kb % a.cin_chunks;
            const int tap = kb / a.cin_chunks;
            const int kw = tap % a.KW, kh = tap / a.KW;
            ptx::tma_load_4d(dA, &tmapA, &full[stage], cc * BLOCK_K, wo0 * a.stride + kw - a.pad_w,
                             ho0 * a.stride + kh - a.pad_h, n0);
            ptx::tma_load_2d(dB, &tmapB, &full[stage], kb * BLOCK_K, nt * a.N_tile);
          }
        }
        __syncwarp();
        if (++stage == a.stages) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (whole warp loops; tcgen05 under elect.sync) =====================
    // pair mode: only the leader CTA issues; its M = 256 instruction reads both CTAs' stages and writes
    // both CTAs' accumulators, and its commits arrive on the barriers of BOTH CTAs.
    const bool pair = CL == 2;
    const uint32_t idesc = ptx::make_idesc_bf16(pair ? 2 * BLOCK_M : BLOCK_M, a.N_tile);
    const uint32_t hi = a.is_stem ? ptx::sw64_desc_hi() : ptx::sw128_desc_hi();
    int stage = 0;
    uint32_t phase = 0;
    int acc = 0;
    uint32_t acc_phase = 0;
    if (!pair || crank == 0)
    for (int it = 0;; ++it) {
      int nt_u, mt_u;
      if (!unit_of_iter(it, nt_u, mt_u)) break;
      ptx::mbar_wait(&tempty[acc], acc_phase ^ 1);
      ptx::tc_fence_after();
      const uint32_t d_tmem = tmem_base + (uint32_t)(acc * 256);
      const int kb0 = nt_u >= a.center_nt ? a.center_kb0 : 0, kb1 = nt_u >= a.center_nt ? kb0 + a.cin_chunks : num_kb;
      for (int kb = kb0; kb < kb1; ++kb) {
        ptx::mbar_wait(&full[stage], phase);
        ptx::tc_fence_after();
        const uint32_t a_lo = ptx::desc_lo(ptx::smem_u32(sA + (size_t)stage * A_BYTES));
        const uint32_t b_lo = ptx::desc_lo(ptx::smem_u32(sB + (size_t)stage * B_BYTES));
        if (ptx::elect_one()) {
          // advance 16 bf16 = 32 bytes along K inside the swizzle row: +2 in 16-byte units
          if (pair) {
            ptx::umma_bf16_lohi_2sm(d_tmem, a_lo, b_lo, hi, idesc, (uint32_t)(kb != kb0));
            ptx::umma_bf16_lohi_2sm(d_tmem, a_lo + 2, b_lo + 2, hi, idesc, 1u);
            ptx::umma_bf16_lohi_2sm(d_tmem, a_lo + 4, b_lo + 4, hi, idesc, 1u);
            ptx::umma_bf16_lohi_2sm(d_tmem, a_lo + 6, b_lo + 6, hi, idesc, 1u);
            ptx::umma_commit_2sm(&empty[stage], (uint16_t)3);      // frees the slot in both CTAs
            if (kb == kb1 - 1) ptx::umma_commit_2sm(&tfull[acc], (uint16_t)3);
          } else {
            ptx::umma_bf16_lohi(d_tmem, a_lo, b_lo, hi, idesc, (uint32_t)(kb != kb0));
            ptx::umma_bf16_lohi(d_tmem, a_lo + 2, b_lo + 2, hi, idesc, 1u);
            if (!a.is_stem) {
              ptx::umma_bf16_lohi(d_tmem, a_lo + 4, b_lo + 4, hi, idesc, 1u);
              ptx::umma_bf16_lohi(d_tmem, a_lo + 6, b_lo + 6, hi, idesc, 1u);
            }
            ptx::umma_commit(&empty[stage]);   // frees the smem slot once these MMAs retire
            if (kb == kb1 - 1) ptx::umma_commit(&tfull[acc]);     // accumulator ready for the epilogue
          }
        }
        __syncwarp();
        if (++stage == a.stages) {
          stage = 0;
          phase ^= 1;
        }
      }
      if (++acc == 2) {
        acc = 0;
        acc_phase ^= 1;
      }
    }
  } else if (a.tma_epi) {
    // ===================== epilogue, two groups of 4 warps (2-5, 6-9) =====================
    // group g takes the 64-column chunks j = g, g+2, ... of every tile (or, when the tile has a single
    // chunk, every other tile: accumulator stage g); each chunk is staged swizzled in the group's 16 KB
    // buffer and written by ONE 4-D TMA store (box = the tile's pixel box x 64 channels; partial
    // tiles are clipped by the tensor map).  The residual comes straight from global memory, loaded
    // one chunk ahead (and before the wait on the accumulator) into registers.
    const int ew = warp - 2, g = ew >> 2;
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const int gt = (ew & 3) * 32 + lane;
    const bool leader = gt == 0;
    for (int i = (int)threadIdx.x - 64; i < a.Cout; i += 256) sBias[i] = a.bias[i];
    ptx::named_bar_sync(3, 256);
    const int cn_full = a.N_tile >> 6;                 // 64-column chunks
    const int cn = cn_full + ((a.N_tile & 32) ? 1 : 0);  // + one 32-column tail chunk
    const int jstep = cn > 1 ? 2 : 1;
    const uint32_t myOut = ptx::smem_u32(sOut + (size_t)g * 16384);
    const int w_l = row % a.bw;
    const int h_l = (row / a.bw) % a.bh;
    const int n_l = row / (a.bw * a.bh);
    int acc = 0, k = 0;
    uint32_t acc_phase = 0, res_phase = 0;
    for (;; ++k) {
      int nt, mt;
      if (!unit_of_iter(k, nt, mt)) break;
      const bool part = cn > 1 ? (g < cn) : ((k & 1) == g);
      if (part) {
        const int tw = mt % a.tiles_w;
        mt /= a.tiles_w;
        const int th = mt % a.tiles_h;
        const int tn = mt / a.tiles_h;
        const int wo = tw * a.bw + w_l, ho = th * a.bh + h_l, n = tn * a.bn + n_l;
        const bool do_res = a.res_C && (wo < a.Wo) && (ho < a.Ho) && (n < a.B);
        const size_t rpix = ((size_t)n * a.res_Hp + ho + a.res_pad) * a.res_Wp + wo + a.res_pad;
        const uint4* rbase = reinterpret_cast<const uint4*>(a.residual + rpix * a.res_C + nt * a.N_tile);
        int j = cn > 1 ? g : 0;
        if (a.res_smem) {
          // Residual through shared memory (long-K convs, where the exposed TMA latency per chunk hides under the
          // tile's MMA time): the leader loads the residual box of the chunk into the staging buffer as soon as the
          // previous store has read it, every thread adds its own row in place (see epi_chunk64_smemres).
          auto issue_res = [&](int jj) {
            if (leader) {
              ptx::bulk_wait_read<0>();
              ptx::mbar_arrive_expect_tx(&resFull[g], 16384u);
              ptx::tma_load_4d(sOut + (size_t)g * 16384, &tmapR, &resFull[g], nt * a.N_tile + 64 * jj, tw * a.bw,
                               th * a.bh, tn * a.bn);
            }
          };
          issue_res(j);                                   // before the wait on the accumulator
          ptx::mbar_wait(&tfull[acc], acc_phase);
          ptx::tc_fence_after();
          for (; j < cn; j += jstep) {
            const bool last = j + jstep >= cn;
            ptx::mbar_wait(&resFull[g], res_phase);
            res_phase ^= 1;
            const uint32_t t_addr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * 256 + 64 * j);
            epi_chunk64_smemres(t_addr, row, myOut, sBias + nt * a.N_tile + 64 * j, (nt * a.N_tile + 64 * j) < a.relu_cols);
            if (last) {
              ptx::tc_fence_before();
              __syncwarp();
              if (lane == 0) {
                if (CL == 2) ptx::mbar_arrive_cluster(&tempty[acc], 0u); else ptx::mbar_arrive(&tempty[acc]);
              }
            }
            ptx::fence_proxy_async();
            ptx::named_bar_sync(1 + g, 128);              // chunk staged
            if (leader) {
              ptx::tma_store_4d(&tmapO, sOut + (size_t)g * 16384, a.out_c_off + nt * a.N_tile + 64 * j, tw * a.bw,
                                th * a.bh, tn * a.bn);
              ptx::bulk_commit();
            }
            if (!last) issue_res(j + jstep);
          }
          if (++acc == 2) {
            acc = 0;
            acc_phase ^= 1;
          }
          continue;
        }
        uint4 rv[8], rn[8];
        if (do_res) {
          const int nv = j < cn_full ? 8 : 4;
#pragma unroll
          for (int i = 0; i < 8; ++i)
            if (i < nv) rv[i] = __ldg(rbase + 8 * j + i);
        }
        ptx::mbar_wait(&tfull[acc], acc_phase);
        ptx::tc_fence_after();
        for (; j < cn; j += jstep) {
          const bool last = j + jstep >= cn;
          const bool full = j < cn_full;
          if (do_res && !last) {
            const int nv = (j + jstep) < cn_full ? 8 : 4;
#pragma unroll
            for (int i = 0; i < 8; ++i)
              if (i < nv) rn[i] = __ldg(rbase + 8 * (j + jstep) + i);
          }
          if (leader) ptx::bulk_wait_read<0>();
          ptx::named_bar_sync(1 + g, 128);              // the group's staging buffer is free again
          const uint32_t t_addr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * 256 + 64 * j);
          if (full)
            epi_chunk64_regs(t_addr, row, myOut, do_res, rv, sBias + nt * a.N_tile + 64 * j, (nt * a.N_tile + 64 * j) < a.relu_cols);
          else
            epi_chunk32_regs(t_addr, row, myOut, do_res, rv, sBias + nt * a.N_tile + 64 * j, (nt * a.N_tile + 64 * j) < a.relu_cols);
          if (last) {
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) {
              if (CL == 2) ptx::mbar_arrive_cluster(&tempty[acc], 0u); else ptx::mbar_arrive(&tempty[acc]);
            }
          }
          ptx::fence_proxy_async();
          ptx::named_bar_sync(1 + g, 128);              // chunk staged
          if (leader) {
            ptx::tma_store_4d(full ? &tmapO : &tmapO32, sOut + (size_t)g * 16384, a.out_c_off + nt * a.N_tile + 64 * j,
                              tw * a.bw, th * a.bh, tn * a.bn);
            ptx::bulk_commit();
          }
          if (!last) {
#pragma unroll
            for (int i = 0; i < 8; ++i) rv[i] = rn[i];
          }
        }
      }
      if (++acc == 2) {
        acc = 0;
        acc_phase ^= 1;
      }
    }
    if (leader) ptx::bulk_wait<0>();
  } else if (warp < 6) {
    // ===================== legacy epilogue (warps 2..5): N_tile not a multiple of 64 =====================
    const int q = warp & 3;                    // TMEM lane quarter this warp may access
    const int row = q * 32 + lane;             // row of the 128-row tile == pixel in the box
    const int w_l = row % a.bw;
    const int h_l = (row / a.bw) % a.bh;
    const int n_l = row / (a.bw * a.bh);
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int it = 0;; ++it) {
      int nt, mt;
      if (!unit_of_iter(it, nt, mt)) break;
      const int tw = mt % a.tiles_w;
      mt /= a.tiles_w;
      const int th = mt % a.tiles_h;
      const int tn = mt / a.tiles_h;
      const int wo = tw * a.bw + w_l, ho = th * a.bh + h_l, n = tn * a.bn + n_l;
      const bool valid = (wo < a.Wo) && (ho < a.Ho) && (n < a.B);
      const size_t pix = ((size_t)n * a.out_Hp + ho + a.out_pad) * a.out_Wp + wo + a.out_pad;
      const size_t rpix = ((size_t)n * a.res_Hp + ho + a.res_pad) * a.res_Wp + wo + a.res_pad;
      ptx::mbar_wait(&tfull[acc], acc_phase);
      ptx::tc_fence_after();
      const uint32_t t_addr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * 256);
      for (int c0 = 0; c0 < a.N_tile; c0 += 32) {
        uint32_t r[32];
        ptx::tmem_ld_32x32b_x32(t_addr + (uint32_t)c0, r);
        ptx::tmem_ld_wait();
        if (valid) {
          const int ch = nt * a.N_tile + c0;
          const float4* b4 = reinterpret_cast<const float4*>(a.bias + ch);
          float v[32];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            float4 b = __ldg(b4 + i);
            v[4 * i + 0] = __uint_as_float(r[4 * i + 0]) + b.x;
            v[4 * i + 1] = __uint_as_float(r[4 * i + 1]) + b.y;
            v[4 * i + 2] = __uint_as_float(r[4 * i + 2]) + b.z;
            v[4 * i + 3] = __uint_as_float(r[4 * i + 3]) + b.w;
          }
          if (a.res_C) {
            const uint4* rp = reinterpret_cast<const uint4*>(a.residual + rpix * a.res_C + ch);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              uint4 rv = __ldg(rp + i);
              v[8 * i + 0] += bf16_lo(rv.x); v[8 * i + 1] += bf16_hi(rv.x);
              v[8 * i + 2] += bf16_lo(rv.y); v[8 * i + 3] += bf16_hi(rv.y);
              v[8 * i + 4] += bf16_lo(rv.z); v[8 * i + 5] += bf16_hi(rv.z);
              v[8 * i + 6] += bf16_lo(rv.w); v[8 * i + 7] += bf16_hi(rv.w);
            }
          }
          if (a.relu) {
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = fmaxf(v[i], 0.0f);
          }
          uint4* op = reinterpret_cast<uint4*>(a.out + pix * a.out_C + a.out_c_off + ch);
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            uint4 o;
            o.x = pack_bf16(v[8 * i + 0], v[8 * i + 1]);
            o.y = pack_bf16(v[8 * i + 2], v[8 * i + 3]);
            o.z = pack_bf16(v[8 * i + 4], v[8 * i + 5]);
            o.w = pack_bf16(v[8 * i + 6], v[8 * i + 7]);
            op[i] = o;
          }
        }
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (CL == 2) ptx::mbar_arrive_cluster(&tempty[acc], 0u); else ptx::mbar_arrive(&tempty[acc]);
      }
      if (++acc == 2) {
        acc = 0;
        acc_phase ^= 1;
      }
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (CL == 2) ptx::cluster_sync();   // nobody leaves while the peer may still multicast / arrive here
  if (warp == 1) {
    __syncwarp();
    ptx::tc_fence_after();
    if (CL == 2) ptx::tmem_dealloc_2sm(tmem_base, 512); else ptx::tmem_dealloc(tmem_base, 512);
  }
}

// no alignment slack: the dynamic shared window of a kernel without static __shared__ starts 1024-aligned
// (the kernel traps if it ever does not); 4 stages of 48 KB + staging only just fit in 227 KB
static inline size_t smem_fixed(int tma_epi, int Cout) {
  return (tma_epi ? 2 * 16384 + (size_t)((Cout + 31) & ~31) * 4 : 0) + 256;
}
static inline size_t stage_bytes(int N_tile, int cluster) {
  return A_BYTES + (size_t)(cluster == 2 ? N_tile / 2 : N_tile) * BLOCK_K * 2;
}
static inline size_t smem_bytes(int N_tile, int stages, int tma_epi, int Cout, int cluster) {
  return smem_fixed(tma_epi, Cout) + (size_t)stages * stage_bytes(N_tile, cluster);
}

}  // namespace convk
