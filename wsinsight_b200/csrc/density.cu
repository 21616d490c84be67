// K5 (BASELINE config 5): per-class density grid of classified objects (patches or cells).
//
// The reference has no heat-map code; the only upstream primitive is the object centre
//   center = np.rint(minx + width / 2)            wsinsight/insightlib/insight_helpers.py:26-27
// so the grid is DEFINED here (parity unpinned upstream, pinned by oracle/density.py): object i with class
// argmax_k prob[i][k] (first maximum, numpy semantics) adds 1 to grid[class][center_y / bin][center_x / bin].
// Ranks bin their own shard; the grids are summed with one NCCL all-reduce (wsinsight_b200/insightlib/density.py).
#include "common.cuh"
#include "../../include/wsi_b200.h"

namespace {

// np.rint(a + w / 2) for integers a, w (float64 arithmetic is exact here): round half to even
__device__ __forceinline__ long long rint_half(long long a, long long w) {
  const long long s = 2 * a + w;
  long long m = s >> 1;                       // floor(s / 2) (arithmetic shift)
  if (s & 1) m += (m & 1);                    // x.5: up if floor is odd
  return m;
}

__global__ void k_density_accumulate(const int32_t* __restrict__ xywh, const float* __restrict__ probs, int n, int K,
                                     int bin_px, int Gx, int Gy, int* __restrict__ grid, int* __restrict__ dropped) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  long long cell = -1;
  if (i < n) {
    const int4 r = __ldg(reinterpret_cast<const int4*>(xywh) + i);
    const long long cx = rint_half(r.x, r.z), cy = rint_half(r.y, r.w);
    int best = 0;
    if (K > 1) {
      float bv = probs[(size_t)i * K];
      for (int k = 1; k < K; ++k) {
        const float v = probs[(size_t)i * K + k];
        if (v > bv) { bv = v; best = k; }
      }
    } else {
      best = 0;                                // K == 1: a single (positive-class) grid counts every object
    }
    const long long gx = cx >= 0 ? cx / bin_px : -1, gy = cy >= 0 ? cy / bin_px : -1;
    if (gx >= 0 && gx < Gx && gy >= 0 && gy < Gy) cell = ((long long)best * Gy + gy) * Gx + gx;
    else atomicAdd(dropped, 1);
  }
  // warp-aggregated atomics: neighbouring rows (x-major sorted patch lists) mostly share a bin
  const unsigned active = __ballot_sync(0xffffffffu, cell >= 0);
  if (cell >= 0) {
    const unsigned peers = __match_any_sync(active, cell);
    if ((int)(__ffs(peers) - 1) == (int)(threadIdx.x & 31)) atomicAdd(grid + cell, __popc(peers));
  }
}

}  // namespace

extern "C" {

// grid int32 [K][Gy][Gx] is ACCUMULATED into (zero it first); d_dropped (optional, int32) counts objects whose
// centre falls outside the grid.  xywh int32 [n][4] (minx, miny, width, height), probs f32 [n][K].
int wsi_density_accumulate(const int32_t* d_xywh, const float* d_probs, int n, int K, int bin_px, int Gx, int Gy,
                           int32_t* d_grid, int32_t* d_dropped, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  WSI_ARG_CHECK(d_xywh && d_probs && d_grid && d_dropped, "wsi_density_accumulate: null pointer");
  WSI_ARG_CHECK(n >= 0 && K >= 1 && bin_px >= 1 && Gx >= 1 && Gy >= 1, "wsi_density_accumulate: bad sizes");
  WSI_ARG_CHECK(((uintptr_t)d_xywh & 15) == 0, "wsi_density_accumulate: xywh must be 16-byte aligned");
  if (n == 0) return WSI_OK;
  k_density_accumulate<<<(n + 255) / 256, 256, 0, stream>>>(d_xywh, d_probs, n, K, bin_px, Gx, Gy, d_grid, d_dropped);
  wsi_count_launch(1);
  WSI_CUDA_CHECK(cudaGetLastError());
  return WSI_OK;
}

}  // extern "C"
