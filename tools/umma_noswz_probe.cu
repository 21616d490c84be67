// Standalone probe: tcgen05.mma with a K-major NO-SWIZZLE A operand whose core matrices OVERLAP.
// A dense row of 16-byte "pixel pairs" sits in shared memory; operand row r is the 32-byte window
// starting at pair (start + r): A(r, k) = dense[(start + r) * 8 + k], k < 16.  In the canonical
// no-swizzle layout a core matrix is 8 rows x 16 B at a 16-byte row pitch, so this is the descriptor
// {start, K-direction core stride = 16 B, M-direction 8-row-group stride = 128 B}.  If the hardware
// accepts it, a 7x7/2 stem can read its overlapping input windows straight from an unexpanded image row.
// The probe tries both assignments of the two strides to the LBO / SBO descriptor fields.
#include "../wsinsight_b200/csrc/common.cuh"
#include <cstdlib>
#include <vector>
void wsi_set_error(const char*, ...) {}
void wsi_count_launch(int) {}

struct PArgs {
  const __nv_bfloat16* x;   // dense [R*8]
  const __nv_bfloat16* w;   // canonical no-swizzle B block: 64 n x 16 k
  int R, start;
  int a_lbo, a_sbo;         // bytes
  int b_lbo, b_sbo;
  float* out;               // [128][64]
};

__global__ void __launch_bounds__(128, 1) k_probe(PArgs a) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sA = smem;
  uint8_t* sB = smem + 32 * 1024;
  __shared__ __align__(8) uint64_t bar_done;
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < a.R * 8; i += 128) reinterpret_cast<__nv_bfloat16*>(sA)[i] = a.x[i];
  for (int i = threadIdx.x; i < 64 * 16; i += 128) reinterpret_cast<__nv_bfloat16*>(sB)[i] = a.w[i];
  if (threadIdx.x == 0) {
    ptx::mbar_init(&bar_done, 1);
    ptx::fence_barrier_init();
  }
  if (warp == 0) {
    ptx::tmem_alloc(&tmem_slot, 64);
    ptx::tmem_relinquish();
  }
  ptx::fence_proxy_async();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = tmem_slot;
  if (warp == 0) {
    if (ptx::elect_one()) {
      const uint32_t idesc = ptx::make_idesc_bf16(128, 64);
      const uint32_t aA = ptx::smem_u32(sA) + (uint32_t)(a.start * 16);
      const uint32_t aB = ptx::smem_u32(sB);
      uint64_t dA = 0, dB = 0;
      dA |= (uint64_t)((aA & 0x3FFFF) >> 4);
      dA |= (uint64_t)(a.a_lbo >> 4) << 16;
      dA |= (uint64_t)(a.a_sbo >> 4) << 32;
      dA |= (uint64_t)1 << 46;
      dB |= (uint64_t)((aB & 0x3FFFF) >> 4);
      dB |= (uint64_t)(a.b_lbo >> 4) << 16;
      dB |= (uint64_t)(a.b_sbo >> 4) << 32;
      dB |= (uint64_t)1 << 46;
      ptx::umma_bf16(tmem, dA, dB, idesc, 0);
      ptx::umma_commit(&bar_done);
    }
    __syncwarp();
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

int main() {
  const int R = 512;
  std::vector<float> fx(R * 8), fw(64 * 16);
  std::vector<__nv_bfloat16> hx(R * 8), hw(64 * 16);
  uint32_t s = 777u;
  for (size_t i = 0; i < fx.size(); ++i) {
    s = s * 1664525u + 1013904223u;
    fx[i] = (float)((int)((s >> 24) & 15) - 8);
    hx[i] = __float2bfloat16(fx[i]);
  }
  // logical weights w[n][k]; stored canonical: (n/8)*256 + (k/8)*128 + (n%8)*16 + (k%8)*2 bytes
  std::vector<float> wl(64 * 16);
  for (int n = 0; n < 64; ++n)
    for (int k = 0; k < 16; ++k) {
      s = s * 1664525u + 1013904223u;
      float v = (float)((int)((s >> 24) & 7) - 4);
      wl[n * 16 + k] = v;
      size_t off = (size_t)(n / 8) * 128 + (size_t)(k / 8) * 64 + (size_t)(n % 8) * 8 + (k % 8);   // elements
      hw[off] = __float2bfloat16(v);
    }
  __nv_bfloat16 *dx, *dw;
  float* d_out;
  cudaMalloc(&dx, hx.size() * 2);
  cudaMalloc(&dw, hw.size() * 2);
  cudaMalloc(&d_out, 128 * 64 * 4);
  cudaMemcpy(dx, hx.data(), hx.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dw, hw.data(), hw.size() * 2, cudaMemcpyHostToDevice);
  cudaFuncSetAttribute(k_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
  // {start, a_lbo, a_sbo, b_lbo, b_sbo}
  const int V[][5] = {
      {0, 16, 128, 128, 256}, {0, 128, 16, 256, 128}, {0, 16, 128, 256, 128}, {0, 128, 16, 128, 256},
      {5, 16, 128, 128, 256}, {5, 128, 16, 256, 128}, {115, 16, 128, 128, 256}, {115, 128, 16, 256, 128},
  };
  for (auto& v : V) {
    PArgs a;
    a.x = dx; a.w = dw; a.R = R; a.start = v[0]; a.a_lbo = v[1]; a.a_sbo = v[2]; a.b_lbo = v[3]; a.b_sbo = v[4];
    a.out = d_out;
    cudaMemset(d_out, 0, 128 * 64 * 4);
    k_probe<<<1, 128, 64 * 1024>>>(a);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) {
      printf("start %d: CUDA error %s\n", v[0], cudaGetErrorString(e));
      return 2;
    }
    std::vector<float> out(128 * 64);
    cudaMemcpy(out.data(), d_out, out.size() * 4, cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int r = 0; r < 128; ++r)
      for (int n = 0; n < 64; ++n) {
        float acc = 0;
        for (int k = 0; k < 16; ++k) acc += fx[(size_t)(v[0] + r) * 8 + k] * wl[n * 16 + k];
        bad += acc != out[r * 64 + n];
      }
    printf("start %3d A{lbo %3d sbo %3d} B{lbo %3d sbo %3d}: %s (%d bad of 8192)\n", v[0], v[1], v[2], v[3], v[4],
           bad ? "MISMATCH" : "ok", bad);
  }
  return 0;
}
