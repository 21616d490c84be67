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
constexpr int THREADS = 192;
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
  int out_pad, res_pad;                      // one-pixel zero border of the destination / residual tensor (0|1)
  int relu;
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

__global__ void __launch_bounds__(THREADS, 1)
    k_conv_igemm(const __grid_constant__ CUtensorMap tmapA, const __grid_constant__ CUtensorMap tmapB,
                 const Args a) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // carve: [stages x A 16KB][stages x B N_tile*128][barriers]
  const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem = smem_raw + (smem_base - ptx::smem_u32(smem_raw));
  const int B_BYTES = a.N_tile * BLOCK_K * 2;           // stage stride (the stem uses half of each stage)
  const int kbytes = a.is_stem ? 64 : 128;              // bytes of K per smem row
  uint8_t* sA = smem;
  uint8_t* sB = smem + (size_t)a.stages * A_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sB + (size_t)a.stages * B_BYTES);
  uint64_t* full = bars;
  uint64_t* empty = bars + a.stages;
  uint64_t* tfull = bars + 2 * a.stages;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m_tiles = a.tiles_w * a.tiles_h * a.tiles_n;
  const int total_tiles = m_tiles * a.tiles_N;
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
      ptx::mbar_init(&tempty[s], 4);
    }
    ptx::fence_barrier_init();
  }
  if (warp == 1) {
    ptx::tmem_alloc(tmem_slot, 512);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const int nt = tile % a.tiles_N;
        int mt = tile / a.tiles_N;
        const int tw = mt % a.tiles_w;
        mt /= a.tiles_w;
        const int th = mt % a.tiles_h;
        const int tn = mt / a.tiles_h;
        const int wo0 = tw * a.bw, ho0 = th * a.bh, n0 = tn * a.bn;
        for (int kb = 0; kb < num_kb; ++kb) {
          ptx::mbar_wait(&empty[stage], phase ^ 1);
          ptx::mbar_arrive_expect_tx(&full[stage], (uint32_t)((BLOCK_M + a.N_tile) * kbytes));
          void* dA = sA + (size_t)stage * A_BYTES;
          void* dB = sB + (size_t)stage * B_BYTES;
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
          if (++stage == a.stages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      const uint32_t idesc = ptx::make_idesc_bf16(BLOCK_M, a.N_tile);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        ptx::mbar_wait(&tempty[acc], acc_phase ^ 1);
        ptx::tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(acc * 256);
        for (int kb = 0; kb < num_kb; ++kb) {
          ptx::mbar_wait(&full[stage], phase);
          ptx::tc_fence_after();
          const uint32_t aA = ptx::smem_u32(sA + (size_t)stage * A_BYTES);
          const uint32_t aB = ptx::smem_u32(sB + (size_t)stage * B_BYTES);
          const uint64_t dA = a.is_stem ? ptx::make_sw64_desc(aA) : ptx::make_sw128_desc(aA);
          const uint64_t dB = a.is_stem ? ptx::make_sw64_desc(aB) : ptx::make_sw128_desc(aB);
          const int nk = a.is_stem ? 2 : 4;
          for (int k = 0; k < nk; ++k) {
            // advance 16 bf16 = 32 bytes along K inside the swizzle row: +2 in 16-byte units
            ptx::umma_bf16(d_tmem, dA + (uint64_t)(2 * k), dB + (uint64_t)(2 * k), idesc, (kb | k) != 0);
          }
          ptx::umma_commit(&empty[stage]);   // frees the smem slot once these MMAs retire
          if (++stage == a.stages) {
            stage = 0;
            phase ^= 1;
          }
        }
        ptx::umma_commit(&tfull[acc]);       // accumulator ready for the epilogue
        if (++acc == 2) {
          acc = 0;
          acc_phase ^= 1;
        }
      }
    }
  } else {
    // ===================== epilogue (warps 2..5) =====================
    const int q = warp & 3;                    // TMEM lane quarter this warp may access
    const int row = q * 32 + lane;             // row of the 128-row tile == pixel in the box
    const int w_l = row % a.bw;
    const int h_l = (row / a.bw) % a.bh;
    const int n_l = row / (a.bw * a.bh);
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      const int nt = tile % a.tiles_N;
      int mt = tile / a.tiles_N;
      const int tw = mt % a.tiles_w;
      mt /= a.tiles_w;
      const int th = mt % a.tiles_h;
      const int tn = mt / a.tiles_h;
      const int wo = tw * a.bw + w_l, ho = th * a.bh + h_l, n = tn * a.bn + n_l;
      const bool valid = (wo < a.Wo) && (ho < a.Ho) && (n < a.B);
      const size_t pix = ((size_t)n * (a.Ho + 2 * a.out_pad) + ho + a.out_pad) * (a.Wo + 2 * a.out_pad) + wo + a.out_pad;
      const size_t rpix = ((size_t)n * (a.Ho + 2 * a.res_pad) + ho + a.res_pad) * (a.Wo + 2 * a.res_pad) + wo + a.res_pad;
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
      if (lane == 0) ptx::mbar_arrive(&tempty[acc]);
      if (++acc == 2) {
        acc = 0;
        acc_phase ^= 1;
      }
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    __syncwarp();
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, 512);
  }
}

static inline size_t smem_bytes(int N_tile, int stages) {
  return 1024 + (size_t)stages * (A_BYTES + (size_t)N_tile * BLOCK_K * 2) + (2 * stages + 4) * 8 + 16;
}

}  // namespace convk
