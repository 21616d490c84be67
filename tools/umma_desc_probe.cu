// Standalone probe: how does tcgen05.mma address a K-major swizzled A operand whose descriptor
// (a) starts at a 128-byte-row offset that is not 1024-byte aligned, and (b) has a stride byte
// offset (8-row group pitch) that is not the canonical 8 rows?  This decides whether a conv
// kernel may keep ONE halo slab in shared memory and feed the kw-shifted taps by moving the
// descriptor start instead of re-loading the slab with TMA.
//
// For every variant the expected result is computed as if the hardware swizzled on ABSOLUTE
// shared-memory address bits (what TMA does when it writes the slab):
//   D[r][n] = sum_k X[start_row + (r/8)*sbo_rows + r%8][k] * B[n][k]
// usage: umma_desc_probe            (runs the built-in variant list, prints one line each)
#include "../wsinsight_b200/csrc/common.cuh"
#include <cstdlib>
#include <vector>
void wsi_set_error(const char*, ...) {}
void wsi_count_launch(int) {}

struct PArgs {
  int rows_box;      // rows of X staged in smem
  int row_bytes;     // 128 (SWIZZLE_128B) or 64 (SWIZZLE_64B)
  int start_row;     // descriptor start = sA + start_row * row_bytes
  int sbo_bytes;     // stride byte offset between 8-row groups
  int base_off;      // descriptor bits 49-51
  float* out;        // [128][64]
};

__global__ void __launch_bounds__(128, 1)
    k_probe(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, PArgs a) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem = smem_raw + (base - ptx::smem_u32(smem_raw));
  uint8_t* sA = smem;
  uint8_t* sB = smem + 64 * 1024;
  __shared__ __align__(8) uint64_t bar_full, bar_done;
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    ptx::mbar_init(&bar_full, 1);
    ptx::mbar_init(&bar_done, 1);
    ptx::fence_barrier_init();
  }
  if (warp == 0) {
    ptx::tmem_alloc(&tmem_slot, 64);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = tmem_slot;
  if (threadIdx.x == 0) {
    ptx::mbar_arrive_expect_tx(&bar_full, (uint32_t)(a.rows_box * a.row_bytes + 64 * a.row_bytes));
    ptx::tma_load_2d(sA, &tmA, &bar_full, 0, 0);
    ptx::tma_load_2d(sB, &tmB, &bar_full, 0, 0);
    ptx::mbar_wait(&bar_full, 0);
    ptx::tc_fence_after();
    const uint32_t idesc = ptx::make_idesc_bf16(128, 64);
    const uint32_t aA = ptx::smem_u32(sA) + (uint32_t)(a.start_row * a.row_bytes);
    const uint32_t aB = ptx::smem_u32(sB);
    uint64_t dA = 0, dB = 0;
    const uint64_t swz = a.row_bytes == 128 ? 2 : 4;
    dA |= (uint64_t)((aA & 0x3FFFF) >> 4);
    dA |= (uint64_t)1 << 16;
    dA |= (uint64_t)(a.sbo_bytes >> 4) << 32;
    dA |= (uint64_t)1 << 46;
    dA |= (uint64_t)(a.base_off & 7) << 49;
    dA |= swz << 61;
    dB |= (uint64_t)((aB & 0x3FFFF) >> 4);
    dB |= (uint64_t)1 << 16;
    dB |= (uint64_t)((8 * a.row_bytes) >> 4) << 32;
    dB |= (uint64_t)1 << 46;
    dB |= swz << 61;
    const int nk = a.row_bytes / 32;
    for (int k = 0; k < nk; ++k) ptx::umma_bf16(tmem, dA + (uint64_t)(2 * k), dB + (uint64_t)(2 * k), idesc, k != 0);
    ptx::umma_commit(&bar_done);
  }
  ptx::mbar_wait(&bar_done, 0);
  ptx::tc_fence_after();
  for (int c0 = 0; c0 < 64; c0 += 32) {
    uint32_t r[32];
    ptx::tmem_ld_32x32b_x32(tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0, r);
    ptx::tmem_ld_wait();
    for (int i = 0; i < 32; ++i) a.out[(warp * 32 + lane) * 64 + c0 + i] = __uint_as_float(r[i]);
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem, 64);
  }
}

typedef CUresult (*fn_t)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                         const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                         CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main() {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
  fn_t enc = (fn_t)p;
  const int R = 256;
  cudaFuncSetAttribute(k_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  float* d_out;
  cudaMalloc(&d_out, 128 * 64 * 4);
  // {row_bytes, start_row, sbo_rows, base_off}
  const int V[][4] = {
      {128, 0, 8, 0},  {128, 1, 8, 0},  {128, 1, 8, 1},  {128, 2, 8, 2},  {128, 0, 10, 0}, {128, 1, 10, 0},
      {128, 1, 10, 1}, {128, 2, 10, 0}, {128, 2, 10, 2}, {128, 0, 16, 0}, {128, 1, 16, 0}, {128, 1, 16, 1},
      {128, 2, 16, 2}, {128, 0, 6, 0},  {128, 2, 6, 0},  {128, 0, 12, 0}, {128, 1, 12, 0}, {128, 2, 12, 0},
      {64, 0, 8, 0},   {64, 8, 16, 0},  {64, 0, 16, 0},  {64, 4, 16, 0},  {64, 1, 8, 0},   {64, 1, 8, 1},
      {64, 3, 10, 0},
  };
  for (auto& v : V) {
    const int rb = v[0], start = v[1], sbo_rows = v[2], bo = v[3];
    const int kk = rb / 2;
    std::vector<__nv_bfloat16> hx((size_t)R * kk), hb((size_t)64 * kk);
    std::vector<float> fx(hx.size()), fb(hb.size());
    uint32_t s = 12345u;
    for (size_t i = 0; i < hx.size(); ++i) {
      s = s * 1664525u + 1013904223u;
      fx[i] = (float)((int)((s >> 24) & 15) - 8);
      hx[i] = __float2bfloat16(fx[i]);
    }
    for (size_t i = 0; i < hb.size(); ++i) {
      s = s * 1664525u + 1013904223u;
      fb[i] = (float)((int)((s >> 24) & 7) - 4);
      hb[i] = __float2bfloat16(fb[i]);
    }
    __nv_bfloat16 *dx, *db;
    cudaMalloc(&dx, hx.size() * 2);
    cudaMalloc(&db, hb.size() * 2);
    cudaMemcpy(dx, hx.data(), hx.size() * 2, cudaMemcpyHostToDevice);
    cudaMemcpy(db, hb.data(), hb.size() * 2, cudaMemcpyHostToDevice);
    CUtensorMap tA, tB;
    cuuint64_t gdA[2] = {(cuuint64_t)kk, (cuuint64_t)R}, gs[1] = {(cuuint64_t)rb};
    cuuint32_t bxA[2] = {(cuuint32_t)kk, (cuuint32_t)R}, es[2] = {1, 1};
    cuuint64_t gdB[2] = {(cuuint64_t)kk, 64};
    cuuint32_t bxB[2] = {(cuuint32_t)kk, 64};
    CUtensorMapSwizzle sw = rb == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B;
    CUresult r1 = enc(&tA, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, dx, gdA, gs, bxA, es, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                      CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    CUresult r2 = enc(&tB, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, db, gdB, gs, bxB, es, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                      CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r1 || r2) {
      printf("encode failed %d %d\n", (int)r1, (int)r2);
      return 1;
    }
    PArgs a;
    a.rows_box = R; a.row_bytes = rb; a.start_row = start; a.sbo_bytes = sbo_rows * rb; a.base_off = bo; a.out = d_out;
    cudaMemset(d_out, 0, 128 * 64 * 4);
    k_probe<<<1, 128, 100 * 1024>>>(tA, tB, a);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) {
      printf("row_bytes %d start %d sbo_rows %d base_off %d: CUDA error %s\n", rb, start, sbo_rows, bo,
             cudaGetErrorString(e));
      return 2;
    }
    std::vector<float> out(128 * 64);
    cudaMemcpy(out.data(), d_out, out.size() * 4, cudaMemcpyDeviceToHost);
    int bad = 0, bad_rows = 0;
    for (int r = 0; r < 128; ++r) {
      const int row = start + (r / 8) * sbo_rows + (r % 8);
      int rb_bad = 0;
      for (int n = 0; n < 64; ++n) {
        float acc = 0;
        for (int k = 0; k < kk; ++k) acc += fx[(size_t)row * kk + k] * fb[(size_t)n * kk + k];
        if (acc != out[r * 64 + n]) ++rb_bad;
      }
      bad += rb_bad;
      bad_rows += rb_bad != 0;
    }
    printf("row_bytes %3d start_row %2d sbo_rows %2d base_off %d: %s (%d bad values in %d rows)\n", rb, start, sbo_rows,
           bo, bad ? "MISMATCH" : "ok", bad, bad_rows);
    cudaFree(dx);
    cudaFree(db);
  }
  return 0;
}
