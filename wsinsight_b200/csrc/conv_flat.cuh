// K4a': 3x3 / stride 1 / pad 1 convolution over a zero-PADDED NHWC activation, as a tcgen05
// implicit GEMM whose nine taps are fed from ONE shared-memory slab per tile.
//
// Why: with per-tap TMA im2col boxes (conv.cuh) the input is re-read nine times from L2 and the
// 128-byte-granular 4-D boxes cap the A-operand feed (ncu, profiles/r01_ncu_conv_baseline_summary.txt:
// layer1 moves 5.55 GB L2->SM per launch for a 411 MB input; tensor pipe 19 % active).  Here the
// activation is stored with a one-pixel zero border, [B][H+2][W+2][C], so that in FLAT pixel
// index p = (n*(H+2) + h)*(W+2) + w the input of tap (kh,kw) for output p is simply row
// p + (kh-1)*(W+2) + (kw-1): a pure row shift.  One contiguous 2-D TMA load brings rows
// [q0 - (W+2) - 1, q0 + T*128 + (W+2) + 1) of a 64-channel chunk into a slab; every tap is an
// MMA whose A descriptor START ADDRESS is moved by (kh*(W+2) + kw) rows.  tcgen05 applies the
// 128B swizzle on absolute shared-memory address bits, so any 128-byte-row start works
// (tools/umma_desc_probe.cu, profiles/r01_umma_desc_probe.txt).  Outputs are computed for every
// padded position; border positions are simply not stored (they stay zero = the next conv's padding).
//
// T = 2 M-tiles (256 output positions) share each weight k-block, halving the B-operand feed;
// small weight sets (<= 80 KB: ResNet layer1) stay RESIDENT in shared memory for the whole kernel.
//
// Warp roles as conv.cuh: warp 0 TMA producer, warp 1 MMA issuer, warps 2-5 and 6-9 two epilogue groups
// (group g owns M-tile g of every super-tile: with N = 64 one group could not drain 128 x 64 accumulators,
// add the residual and stage the store in the ~1800 cycles the tensor pipe needs for the next tile --
// ncu: tensor pipe 39 % / 55 % active with / without residual on ResNet34 layer1).
#pragma once
#include "conv.cuh"

namespace convf {

constexpr int THREADS = 320;   // warp 0 producer, warp 1 MMA, warps 2-5 / 6-9: epilogue group of M-tile 0 / 1

struct Args {
  int Hp, Wp, H, W, B;          // stored geometry shared by input, output and residual; H x W = OUTPUT interior
  int pad;                      // border of the OUTPUT interior inside the stored geometry (a 'valid' 3x3 conv is the
                                // same centred flat conv whose output border is one pixel wider than its input's)
  int cluster;                  // 1, or 2: CTA pairs (k_conv_flat<2>, launched as 2-CTA clusters)
  int ksteps;                   // K = 16 MMA steps per tap and 64-channel chunk: 4, or 2 when only 32 channels are real
  long long total_rows;         // B * Hp * Wp flat positions
  int num_tiles;                // super-tiles of T*128 positions
  int T;                        // M-tiles per super-tile (1 or 2)
  int N;                        // Cout (UMMA N, multiple of 16, <= 128)
  int chunks;                   // Cin / 64
  int resident_b;               // all 9*chunks weight k-blocks stay in smem
  int slab_rows;                // rows per slab (multiple of 8)
  int slab_loads, slab_box_rows;
  int stagesA, stagesB;         // slab ring / weight ring (stagesB == 9*chunks when resident)
  int relu, has_res;
  int out_C, out_c_off;         // channels of the destination tensor / concat offset
  int tma_epi;                  // N % 64 == 0: epilogue staged in smem and written by TMA stores
  int res_smem;                 // residual tiles are TMA-loaded into the output staging buffers (needs nOut == 2)
  int res_C, res_c_off;         // channels of the residual tensor (row pitch) and the first one read: `residual` points at it
  int nOut, nRes;               // output staging buffers (16 KB each) PER epilogue group; nRes unused (residual
                                // chunks are prefetched from global memory into registers one chunk ahead)
  const __nv_bfloat16* residual;
  const float* bias;
  __nv_bfloat16* out;
};

// MMA issue loop of the flat kernel, specialised on T (M-tiles per super-tile) and weight residency so
// that the nine taps unroll into straight-line UTCHMMA sequences: the issuing warp, not the tensor pipe,
// was the limiter (ncu: ~975 cycles per tap for 256-512 cycles of MMA work) while the loop carried
// runtime branches, constant reloads and one elect region per tap.
struct MmaCtx {
  uint32_t sA_lo, sB_lo;          // descriptor low words of slab 0 / weight block 0
  uint32_t slab_step, b_step;     // (bytes >> 4) between slabs / weight blocks
  uint32_t row_step;              // Wp * 8: one image row of the padded layout in descriptor units
  uint32_t tmem_base, idesc, hi;
  int N, chunks, stagesA, stagesB, num_tiles;
  int first_unit, unit_step, num_units;   // work units of this CTA (pair mode: one unit = two consecutive super-tiles)
  uint64_t *aFull, *aEmpty, *bFull, *bEmpty, *tfull, *tempty;
};

template <bool PAIR>
__device__ __forceinline__ void umma1(uint32_t d, uint32_t a_lo, uint32_t b_lo, uint32_t hi, uint32_t idesc, uint32_t acc) {
  if (PAIR) ptx::umma_bf16_lohi_2sm(d, a_lo, b_lo, hi, idesc, acc); else ptx::umma_bf16_lohi(d, a_lo, b_lo, hi, idesc, acc);
}

template <int T, int KS, bool PAIR>
__device__ __forceinline__ void issue_tap(uint32_t d_tmem, uint32_t a_lo, uint32_t b_lo, uint32_t hi, uint32_t idesc,
                                          uint32_t accum_first, uint32_t N) {
  umma1<PAIR>(d_tmem, a_lo + 0, b_lo + 0, hi, idesc, accum_first);
  umma1<PAIR>(d_tmem, a_lo + 2, b_lo + 2, hi, idesc, 1u);
  if (KS == 4) {
    umma1<PAIR>(d_tmem, a_lo + 4, b_lo + 4, hi, idesc, 1u);
    umma1<PAIR>(d_tmem, a_lo + 6, b_lo + 6, hi, idesc, 1u);
  }
  if (T == 2) {
    const uint32_t d1 = d_tmem + N, a1 = a_lo + 1024u;   // second M-tile: +128 rows of 128 B (>> 4)
    umma1<PAIR>(d1, a1 + 0, b_lo + 0, hi, idesc, accum_first);
    umma1<PAIR>(d1, a1 + 2, b_lo + 2, hi, idesc, 1u);
    if (KS == 4) {
      umma1<PAIR>(d1, a1 + 4, b_lo + 4, hi, idesc, 1u);
      umma1<PAIR>(d1, a1 + 6, b_lo + 6, hi, idesc, 1u);
    }
  }
}

template <bool PAIR>
__device__ __forceinline__ void commit1(uint64_t* bar) {
  if (PAIR) ptx::umma_commit_2sm(bar, (uint16_t)3); else ptx::umma_commit(bar);   // pair: arrives in both CTAs
}

template <int T, bool RESIDENT, int KS, bool PAIR>
__device__ __forceinline__ void mma_loop(const MmaCtx& m) {
  int sa = 0, sb = 0, acc = 0;
  uint32_t pa = 0, pb = 0, pacc = 0;
  bool first_tile = true;
  for (int u = m.first_unit; u < m.num_units; u += m.unit_step) {
    ptx::mbar_wait(&m.tempty[acc], pacc ^ 1);
    const uint32_t d_tmem = m.tmem_base + (uint32_t)(acc * T) * (uint32_t)m.N;
    for (int c = 0; c < m.chunks; ++c) {
      ptx::mbar_wait(&m.aFull[sa], pa);
      const uint32_t slab_lo = m.sA_lo + (uint32_t)sa * m.slab_step;
      const bool last_c = c == m.chunks - 1;
      if (RESIDENT) {
        if (first_tile)
          for (int tap = 0; tap < 9; ++tap) ptx::mbar_wait(&m.bFull[tap * m.chunks + c], 0);
        ptx::tc_fence_after();
        if (ptx::elect_one()) {
#pragma unroll
          for (int tap = 0; tap < 9; ++tap) {
            const uint32_t a_lo = slab_lo + (uint32_t)(tap / 3) * m.row_step + (uint32_t)(tap % 3) * 8u;
            const uint32_t b_lo = m.sB_lo + (uint32_t)(tap * m.chunks + c) * m.b_step;
            issue_tap<T, KS, PAIR>(d_tmem, a_lo, b_lo, m.hi, m.idesc, tap == 0 ? (uint32_t)(c != 0) : 1u, (uint32_t)m.N);
          }
          commit1<PAIR>(&m.aEmpty[sa]);
          if (last_c) commit1<PAIR>(&m.tfull[acc]);
        }
        __syncwarp();
      } else {
#pragma unroll
        for (int tap = 0; tap < 9; ++tap) {
          ptx::mbar_wait(&m.bFull[sb], pb);
          ptx::tc_fence_after();
          if (ptx::elect_one()) {
            const uint32_t a_lo = slab_lo + (uint32_t)(tap / 3) * m.row_step + (uint32_t)(tap % 3) * 8u;
            const uint32_t b_lo = m.sB_lo + (uint32_t)sb * m.b_step;
            issue_tap<T, KS, PAIR>(d_tmem, a_lo, b_lo, m.hi, m.idesc, tap == 0 ? (uint32_t)(c != 0) : 1u, (uint32_t)m.N);
            commit1<PAIR>(&m.bEmpty[sb]);
            if (tap == 8) {
              commit1<PAIR>(&m.aEmpty[sa]);
              if (last_c) commit1<PAIR>(&m.tfull[acc]);
            }
          }
          __syncwarp();
          if (++sb == m.stagesB) { sb = 0; pb ^= 1; }
        }
      }
      if (++sa == m.stagesA) { sa = 0; pa ^= 1; }
    }
    first_tile = false;
    if (++acc == 2) { acc = 0; pacc ^= 1; }
  }
}

// CL = 1: one CTA per super-tile.  CL = 2: CTA PAIRS (cta_group::2): the leader's tcgen05.mma spans both SMs
// (M = 256: M-tile t of BOTH CTAs' slabs against the weight k-block whose N/2-row halves sit in the two CTAs), so each
// SM reads 4 KB of A + N*16 B of B per K step instead of 4 KB + N*32 B: with N = 64 / 128 the single-CTA kernel is
// bound by shared-memory operand bandwidth (128 B/clk: 67 % / 100 % of the tensor rate before any other smem traffic).
template <int CL>
__global__ void __launch_bounds__(THREADS, 1)
    k_conv_flat(const __grid_constant__ CUtensorMap tmapA, const __grid_constant__ CUtensorMap tmapB,
                const __grid_constant__ CUtensorMap tmapO, const __grid_constant__ CUtensorMap tmapO32,
                const __grid_constant__ CUtensorMap tmapR, const Args a) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem = smem_raw + (smem_base - ptx::smem_u32(smem_raw));
  const int slab_bytes = a.slab_rows * 128;
  const int b_bytes = (CL == 2 ? a.N / 2 : a.N) * 128;      // per-CTA bytes of one weight k-block
  uint8_t* sA = smem;
  uint8_t* sB = smem + (size_t)a.stagesA * slab_bytes;
  uint8_t* sOut = sB + (size_t)a.stagesB * b_bytes;                 // T groups x nOut x 16 KB (1024-aligned)
  float* sBias = reinterpret_cast<float*>(sOut + (size_t)a.T * a.nOut * 16384);
  uint64_t* bars = reinterpret_cast<uint64_t*>(sBias + 128);
  uint64_t* aFull = bars;
  uint64_t* aEmpty = aFull + a.stagesA;
  uint64_t* bFull = aEmpty + a.stagesA;
  uint64_t* bEmpty = bFull + a.stagesB;
  uint64_t* tfull = bEmpty + a.stagesB;
  uint64_t* tempty = tfull + 2;
  uint64_t* resFull = tempty + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(resFull + 4);     // resFull[group * 2 + buffer]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int TM = a.T * 128;
  // work units: one super-tile per CTA and iteration; in pair mode unit u of the PAIR is super-tiles 2u, 2u + 1 and
  // CTA `crank` takes 2u + crank, clamped to the last one (an odd count makes the last pair compute it twice:
  // identical stores)
  const uint32_t crank = CL == 2 ? ptx::cluster_ctarank() : 0u;
  const int num_units = CL == 2 ? (a.num_tiles + 1) / 2 : a.num_tiles;
  const int first_unit = CL == 2 ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int unit_step = CL == 2 ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  auto tile_of = [&](int u) { return CL == 2 ? min(2 * u + (int)crank, a.num_tiles - 1) : u; };

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tmapA);
    ptx::prefetch_tmap(&tmapB);
    for (int s = 0; s < a.stagesA; ++s) {
      ptx::mbar_init(&aFull[s], 1);
      ptx::mbar_init(&aEmpty[s], 1);
    }
    for (int s = 0; s < a.stagesB; ++s) {
      ptx::mbar_init(&bFull[s], 1);
      ptx::mbar_init(&bEmpty[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      ptx::mbar_init(&tfull[s], 1);
      ptx::mbar_init(&tempty[s], (uint32_t)((a.tma_epi ? 4 * a.T : 4) * CL));
      ptx::mbar_init(&resFull[s], 1);
      ptx::mbar_init(&resFull[2 + s], 1);
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
  if (CL == 2) ptx::cluster_sync();           // the peer's barriers are initialised before anything remote happens
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // barriers, tensor memory and descriptors are set up: only now wait for the producing kernel of the stream (its
  // tail overlapped this prologue), and let the consuming kernel do the same with ours
  ptx::pdl_wait();
  ptx::pdl_launch_dependents();

  if (warp == 0) {
    // ===================== TMA producer =====================
    {
      int sa = 0, sb = 0;
      uint32_t pa = 0, pb = 0;
      bool first = true;
      const bool arm = CL == 1 || crank == 0;     // pair mode: only the leader arms, for the bytes of BOTH CTAs
      for (int u = first_unit; u < num_units; u += unit_step) {
        const long long q0 = (long long)tile_of(u) * TM;
        const int row0 = (int)(q0 - a.Wp - 1);
        if (a.has_res && q0 < a.total_rows && ptx::elect_one()) {
          // the epilogue reads this super-tile's residual rows stagesA slabs from now: have them in L2 by then
          // (HBM latency under load exceeds the one-tile lead of the register prefetch)
          const long long rows = a.total_rows - q0 < TM ? a.total_rows - q0 : TM;
          ptx::bulk_prefetch_l2(a.residual - a.res_c_off + (size_t)q0 * a.res_C, (uint32_t)(rows * a.res_C * 2));
        }
        __syncwarp();
        for (int c = 0; c < a.chunks; ++c) {
          ptx::mbar_wait(&aEmpty[sa], pa ^ 1);
          if (ptx::elect_one()) {
            if (arm) ptx::mbar_arrive_expect_tx(&aFull[sa], (uint32_t)(slab_bytes * CL));
            for (int i = 0; i < a.slab_loads; ++i) {
              void* dst = sA + (size_t)sa * slab_bytes + (size_t)i * a.slab_box_rows * 128;
              if (CL == 2)
                ptx::tma_load_2d_2sm(dst, &tmapA, ptx::leader_bar_addr(&aFull[sa]), c * 64, row0 + i * a.slab_box_rows);
              else
                ptx::tma_load_2d(dst, &tmapA, &aFull[sa], c * 64, row0 + i * a.slab_box_rows);
            }
          }
          __syncwarp();
          if (++sa == a.stagesA) { sa = 0; pa ^= 1; }
          if (a.resident_b) {
            if (first && ptx::elect_one())
              for (int tap = 0; tap < 9; ++tap) {
                const int kb = tap * a.chunks + c;
                if (arm) ptx::mbar_arrive_expect_tx(&bFull[kb], (uint32_t)(b_bytes * CL));
                if (CL == 2)
                  ptx::tma_load_2d_2sm(sB + (size_t)kb * b_bytes, &tmapB, ptx::leader_bar_addr(&bFull[kb]), kb * 64,
                                       (int)crank * (a.N >> 1));
                else
                  ptx::tma_load_2d(sB + (size_t)kb * b_bytes, &tmapB, &bFull[kb], kb * 64, 0);
              }
            __syncwarp();
          } else {
            for (int tap = 0; tap < 9; ++tap) {
              const int kb = tap * a.chunks + c;
              ptx::mbar_wait(&bEmpty[sb], pb ^ 1);
              if (ptx::elect_one()) {
                if (arm) ptx::mbar_arrive_expect_tx(&bFull[sb], (uint32_t)(b_bytes * CL));
                if (CL == 2)
                  ptx::tma_load_2d_2sm(sB + (size_t)sb * b_bytes, &tmapB, ptx::leader_bar_addr(&bFull[sb]), kb * 64,
                                       (int)crank * (a.N >> 1));
                else
                  ptx::tma_load_2d(sB + (size_t)sb * b_bytes, &tmapB, &bFull[sb], kb * 64, 0);
              }
              __syncwarp();
              if (++sb == a.stagesB) { sb = 0; pb ^= 1; }
            }
          }
        }
        first = false;
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    // The whole warp runs the loop (warp-uniform control flow); the tcgen05 instructions sit under
    // elect.sync so that ptxas emits them straight-line (an `if (lane == 0)` region makes it wrap
    // every UTCHMMA in an elect/branch loop: ~100 cycles per MMA, 3x the N=64 MMA itself).
    MmaCtx m;
    m.sA_lo = ptx::desc_lo(ptx::smem_u32(sA));
    m.sB_lo = ptx::desc_lo(ptx::smem_u32(sB));
    m.slab_step = (uint32_t)slab_bytes >> 4;
    m.b_step = (uint32_t)b_bytes >> 4;
    m.row_step = (uint32_t)a.Wp * 8u;
    m.tmem_base = tmem_base;
    m.idesc = ptx::make_idesc_bf16(CL == 2 ? 256 : 128, a.N);
    m.hi = ptx::sw128_desc_hi();
    m.N = a.N; m.chunks = a.chunks; m.stagesA = a.stagesA; m.stagesB = a.stagesB; m.num_tiles = a.num_tiles;
    m.first_unit = first_unit; m.unit_step = unit_step; m.num_units = num_units;
    m.aFull = aFull; m.aEmpty = aEmpty; m.bFull = bFull; m.bEmpty = bEmpty; m.tfull = tfull; m.tempty = tempty;
    constexpr bool PAIR = CL == 2;
    if (!PAIR || crank == 0) {                      // pair mode: only the leader CTA issues
      if (a.ksteps == 2) {
        if (a.T == 2) {
          if (a.resident_b) mma_loop<2, true, 2, PAIR>(m); else mma_loop<2, false, 2, PAIR>(m);
        } else {
          if (a.resident_b) mma_loop<1, true, 2, PAIR>(m); else mma_loop<1, false, 2, PAIR>(m);
        }
      } else if (a.T == 2) {
        if (a.resident_b) mma_loop<2, true, 4, PAIR>(m); else mma_loop<2, false, 4, PAIR>(m);
      } else {
        if (a.resident_b) mma_loop<1, true, 4, PAIR>(m); else mma_loop<1, false, 4, PAIR>(m);
      }
    }
  } else {
    // ===================== epilogue (warps 2..5) =====================
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const int img = a.Hp * a.Wp;
    int acc = 0;
    uint32_t pacc = 0;
    if (a.tma_epi) {
      const int g = (warp - 2) >> 2;                    // epilogue group == M-tile of the super-tile
      const int gt = ((warp - 2) & 3) * 32 + lane;      // 0..127 inside the group
      const bool leader = gt == 0;
      const int cn_full = a.N >> 6;                     // 64-column chunks per accumulator tile
      const int cn = cn_full + ((a.N & 32) ? 1 : 0);    // + one 32-column tail chunk (N = 32, 96)
      for (int i = (int)threadIdx.x - 64; i < a.N; i += 256) sBias[i] = a.bias[i];
      ptx::named_bar_sync(3, 256);                      // bias staged (both groups)
      if (g < a.T) {
        uint8_t* myOut = sOut + (size_t)g * a.nOut * 16384;
        int k = 0;
        if (a.res_smem) {
          // Residual through shared memory: the group leader TMA-loads the residual tile of chunk k + 1 into the OTHER
          // staging buffer (after the store issued from it has finished reading), every thread adds its own row in
          // place, and the same buffer is then TMA-stored as the output.
          uint64_t* rf = resFull + g * 2;
          if (leader && first_unit < num_units) {
            ptx::mbar_arrive_expect_tx(&rf[0], 16384u);
            ptx::tma_load_2d(myOut, &tmapR, &rf[0], 0, (int)((long long)tile_of(first_unit) * TM + g * 128));
          }
          for (int u = first_unit; u < num_units; u += unit_step) {
            const long long q0 = (long long)tile_of(u) * TM + g * 128;
            const long long p = q0 + row;
            const int n = (int)(p / img);
            const int rem = (int)(p - (long long)n * img);
            const int hp = rem / a.Wp, wp = rem - hp * a.Wp;
            const bool valid = (n < a.B) && (hp >= a.pad) && (hp < a.pad + a.H) && (wp >= a.pad) && (wp < a.pad + a.W);
            const int un = u + unit_step;
            const long long q0n = un < num_units ? (long long)tile_of(un) * TM + g * 128 : -1;
            ptx::mbar_wait(&tfull[acc], pacc);
            ptx::tc_fence_after();
            for (int j = 0; j < cn_full; ++j, ++k) {
              const bool last = j == cn_full - 1;
              const int ob = k & 1;
              if (leader && (!last || q0n >= 0)) {
                ptx::bulk_wait_read<0>();               // the store issued from the other buffer has read it
                ptx::mbar_arrive_expect_tx(&rf[ob ^ 1], 16384u);
                ptx::tma_load_2d(myOut + (size_t)(ob ^ 1) * 16384, &tmapR, &rf[ob ^ 1], last ? 0 : 64 * (j + 1),
                                 (int)(last ? q0n : q0));
              }
              ptx::mbar_wait(&rf[ob], (uint32_t)((k >> 1) & 1));
              const uint32_t t_addr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)((acc * a.T + g) * a.N + 64 * j);
              convk::epi_chunk64_smemres(t_addr, row, ptx::smem_u32(myOut + (size_t)ob * 16384), sBias + 64 * j, a.relu,
                                         valid);
              if (last) {
                ptx::tc_fence_before();
                __syncwarp();
                if (lane == 0) {
                  if (CL == 2) ptx::mbar_arrive_cluster(&tempty[acc], 0u); else ptx::mbar_arrive(&tempty[acc]);
                }
              }
              ptx::fence_proxy_async();
              ptx::named_bar_sync(1 + g, 128);          // chunk staged
              if (leader) {
                ptx::tma_store_2d(&tmapO, myOut + (size_t)ob * 16384, a.out_c_off + 64 * j, (int)q0);
                ptx::bulk_commit();
              }
            }
            if (++acc == 2) { acc = 0; pacc ^= 1; }
          }
          if (leader) ptx::bulk_wait<0>();
        } else {
        // Residual rows are prefetched into registers one CHUNK ahead, across tiles: the first chunk of the next
        // super-tile is requested while the last chunk of this one is processed.  (Requesting it only at the top of
        // the next iteration put the whole HBM/L2 latency, ~1.5x the MMA time of an N = 64 tile, on the epilogue's
        // critical path: ncu long-scoreboard stalls, 0.30 ms against 0.215 ms without a residual.)
        uint4 rv[8], rn[8];
        const int nv0 = cn_full > 0 ? 8 : 4;
        if (a.has_res && first_unit < num_units) {
          const long long p0 = (long long)tile_of(first_unit) * TM + g * 128 + row;
          if (p0 < a.total_rows) {
            const uint4* rb = reinterpret_cast<const uint4*>(a.residual + (size_t)p0 * a.res_C);
#pragma unroll
            for (int i = 0; i < 8; ++i)
              if (i < nv0) rv[i] = __ldg(rb + i);
          }
        }
        for (int u = first_unit; u < num_units; u += unit_step) {
          const long long q0 = (long long)tile_of(u) * TM + g * 128;
          const long long p = q0 + row;
          const int n = (int)(p / img);
          const int rem = (int)(p - (long long)n * img);
          const int hp = rem / a.Wp, wp = rem - hp * a.Wp;
          const bool valid = (n < a.B) && (hp >= a.pad) && (hp < a.pad + a.H) && (wp >= a.pad) && (wp < a.pad + a.W);
          const bool do_res = a.has_res && p < a.total_rows;
          const uint4* rbase = reinterpret_cast<const uint4*>(a.residual + (size_t)p * a.res_C);
          const int un = u + unit_step;
          const long long pn = un < num_units ? (long long)tile_of(un) * TM + g * 128 + row : a.total_rows;
          const uint4* rnext = reinterpret_cast<const uint4*>(a.residual + (size_t)pn * a.res_C);
          ptx::mbar_wait(&tfull[acc], pacc);
          ptx::tc_fence_after();
          for (int j = 0; j < cn; ++j, ++k) {
            const bool last = j == cn - 1;
            const bool full = j < cn_full;
            if (do_res && !last) {
              const int nv = (j + 1) < cn_full ? 8 : 4;
#pragma unroll
              for (int i = 0; i < 8; ++i)
                if (i < nv) rn[i] = __ldg(rbase + 8 * (j + 1) + i);
            }
            if (a.has_res && last && pn < a.total_rows) {
#pragma unroll
              for (int i = 0; i < 8; ++i)
                if (i < nv0) rn[i] = __ldg(rnext + i);
            }
            const int ob = a.nOut == 2 ? (k & 1) : 0;
            if (leader) {
              if (a.nOut == 2) ptx::bulk_wait_read<1>(); else ptx::bulk_wait_read<0>();
            }
            ptx::named_bar_sync(1 + g, 128);            // the group's staging buffer ob is free again
            const uint32_t t_addr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)((acc * a.T + g) * a.N + 64 * j);
            if (full)
              convk::epi_chunk64_regs(t_addr, row, ptx::smem_u32(myOut + (size_t)ob * 16384), do_res, rv, sBias + 64 * j,
                                      a.relu, valid);
            else
              convk::epi_chunk32_regs(t_addr, row, ptx::smem_u32(myOut + (size_t)ob * 16384), do_res, rv, sBias + 64 * j,
                                      a.relu, valid);
            if (last) {                                 // this group's accumulator tile fully read: hand it back
              ptx::tc_fence_before();
              __syncwarp();
              if (lane == 0) {
                if (CL == 2) ptx::mbar_arrive_cluster(&tempty[acc], 0u); else ptx::mbar_arrive(&tempty[acc]);
              }
            }
            ptx::fence_proxy_async();
            ptx::named_bar_sync(1 + g, 128);            // chunk staged
            if (leader) {
              ptx::tma_store_2d(full ? &tmapO : &tmapO32, myOut + (size_t)ob * 16384, a.out_c_off + 64 * j, (int)q0);
              ptx::bulk_commit();
            }
            if (a.has_res) {
#pragma unroll
              for (int i = 0; i < 8; ++i) rv[i] = rn[i];
            }
          }
          if (++acc == 2) { acc = 0; pacc ^= 1; }
        }
        if (leader) ptx::bulk_wait<0>();
        }
      }
    } else if (warp < 6)
    for (int u = first_unit; u < num_units; u += unit_step) {
      const long long q0 = (long long)tile_of(u) * TM;
      ptx::mbar_wait(&tfull[acc], pacc);
      ptx::tc_fence_after();
      for (int t = 0; t < a.T; ++t) {
        const long long p = q0 + t * 128 + row;
        const int n = (int)(p / img);
        const int rem = (int)(p - (long long)n * img);
        const int hp = rem / a.Wp, wp = rem - hp * a.Wp;
        const bool valid = (n < a.B) && (hp >= a.pad) && (hp < a.pad + a.H) && (wp >= a.pad) && (wp < a.pad + a.W);
        const uint32_t t_addr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)((acc * a.T + t) * a.N);
        for (int c0 = 0; c0 < a.N; c0 += 32) {
          uint32_t r[32];
          ptx::tmem_ld_32x32b_x32(t_addr + (uint32_t)c0, r);
          ptx::tmem_ld_wait();
          if (valid) {
            const float4* b4 = reinterpret_cast<const float4*>(a.bias + c0);
            float v[32];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              float4 b = __ldg(b4 + i);
              v[4 * i + 0] = __uint_as_float(r[4 * i + 0]) + b.x;
              v[4 * i + 1] = __uint_as_float(r[4 * i + 1]) + b.y;
              v[4 * i + 2] = __uint_as_float(r[4 * i + 2]) + b.z;
              v[4 * i + 3] = __uint_as_float(r[4 * i + 3]) + b.w;
            }
            if (a.has_res) {
              const uint4* rp = reinterpret_cast<const uint4*>(a.residual + (size_t)p * a.res_C + c0);
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                uint4 rv = __ldg(rp + i);
                v[8 * i + 0] += convk::bf16_lo(rv.x); v[8 * i + 1] += convk::bf16_hi(rv.x);
                v[8 * i + 2] += convk::bf16_lo(rv.y); v[8 * i + 3] += convk::bf16_hi(rv.y);
                v[8 * i + 4] += convk::bf16_lo(rv.z); v[8 * i + 5] += convk::bf16_hi(rv.z);
                v[8 * i + 6] += convk::bf16_lo(rv.w); v[8 * i + 7] += convk::bf16_hi(rv.w);
              }
            }
            if (a.relu) {
#pragma unroll
              for (int i = 0; i < 32; ++i) v[i] = fmaxf(v[i], 0.0f);
            }
            uint4* op = reinterpret_cast<uint4*>(a.out + (size_t)p * a.out_C + a.out_c_off + c0);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              uint4 o;
              o.x = convk::pack_bf16(v[8 * i + 0], v[8 * i + 1]);
              o.y = convk::pack_bf16(v[8 * i + 2], v[8 * i + 3]);
              o.z = convk::pack_bf16(v[8 * i + 4], v[8 * i + 5]);
              o.w = convk::pack_bf16(v[8 * i + 6], v[8 * i + 7]);
              op[i] = o;
            }
          }
        }
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (CL == 2) ptx::mbar_arrive_cluster(&tempty[acc], 0u); else ptx::mbar_arrive(&tempty[acc]);
      }
      if (++acc == 2) { acc = 0; pacc ^= 1; }
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (CL == 2) ptx::cluster_sync();   // nobody leaves while the peer may still arrive here
  if (warp == 1) {
    __syncwarp();
    ptx::tc_fence_after();
    if (CL == 2) ptx::tmem_dealloc_2sm(tmem_base, 512); else ptx::tmem_dealloc(tmem_base, 512);
  }
}

static inline size_t smem_bytes(const Args& a) {
  return 1024 + (size_t)a.stagesA * a.slab_rows * 128 + (size_t)a.stagesB * (a.cluster == 2 ? a.N / 2 : a.N) * 128 +
         (size_t)(a.T * a.nOut) * 16384 + 512 + (size_t)(2 * a.stagesA + 2 * a.stagesB + 8) * 8 + 16;
}

}  // namespace convf
