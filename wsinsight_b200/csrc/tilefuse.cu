// a13 / N3: the batch step of the cell-model tile stitcher, fused.
//
// Replaces wsinsight/modellib/tilefuse.py:219-278 TileRemapStitcher.accumulate_batch_torch:
//   softmax(np)[:, 1], softmax(tp)                      (:240-241)
//   F.interpolate(..., (S, S), bilinear, align_corners=False) of np, hv (x alpha), tp   (:244-249)
//   tp / (sum_k tp + 1e-8)                              (:252)
//   D2H of (B, S, S, 3 + K) fp32 and a Python paste loop into host canvases at coords + halo, clipped (:255-278)
// with ONE kernel that writes straight into device-resident canvases: no (B,S,S,3+K) intermediate, no host copy.
// The softmax is evaluated at the four source pixels of every output pixel (the reference's order: softmax on the
// model grid, then interpolation).  Tiles of one call must not overlap (the reference's interiors never do).
#include "common.cuh"
#include "../../include/wsi_b200.h"

namespace {

constexpr int TF_MAXK = 16;

struct TFArgs {
  const float* np_logits;   // [B][2][h][h]
  const float* hv;          // [B][2][h][h]
  const float* tp_logits;   // [B][K][h][h]
  const int32_t* coords;    // [B][cstride] (minx, miny, ...) of the tile in canvas pixels, before the halo offset
  int cstride;
  int B, K, h, S, halo;
  float alpha, scale;       // scale = (float)h / S (PyTorch area_pixel_compute_scale, align_corners = False)
  int W, H;                 // canvas size
  float* np_map;            // [H][W]
  float* hv_map;            // [H][W][2]
  float* tp_map;            // [H][W][K]
};

// PyTorch upsample_bilinear2d source index / weights (align_corners = False)
__device__ __forceinline__ void src_index(float scale, int dst, int size, int& i0, int& i1, float& l0, float& l1) {
  float r = scale * ((float)dst + 0.5f) - 0.5f;
  r = r < 0.f ? 0.f : r;
  i0 = (int)r;
  i1 = i0 + ((i0 < size - 1) ? 1 : 0);
  l1 = r - (float)i0;
  l0 = 1.f - l1;
}

__global__ void __launch_bounds__(256) k_tilefuse(const TFArgs a) {
  const int b = blockIdx.z;
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
  if (x >= a.S || y >= a.S) return;
  const long long cx = (long long)a.coords[(size_t)b * a.cstride] + a.halo + x;
  const long long cy = (long long)a.coords[(size_t)b * a.cstride + 1] + a.halo + y;
  if (cx < 0 || cy < 0 || cx >= a.W || cy >= a.H) return;       // clipped paste (:262-272)
  int x0, x1, y0, y1;
  float wx0, wx1, wy0, wy1;
  src_index(a.scale, x, a.h, x0, x1, wx0, wx1);
  src_index(a.scale, y, a.h, y0, y1, wy0, wy1);
  const int hh = a.h * a.h;
  const int o00 = y0 * a.h + x0, o01 = y0 * a.h + x1, o10 = y1 * a.h + x0, o11 = y1 * a.h + x1;
  auto lerp = [&](float v00, float v01, float v10, float v11) {
    return wy0 * (wx0 * v00 + wx1 * v01) + wy1 * (wx0 * v10 + wx1 * v11);
  };
  const size_t pix = (size_t)cy * a.W + cx;
  // nuclei probability: softmax over 2 channels, channel 1
  {
    const float* p = a.np_logits + (size_t)b * 2 * hh;
    float v[4];
    const int o[4] = {o00, o01, o10, o11};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float l0 = __ldg(p + o[i]), l1 = __ldg(p + hh + o[i]);
      const float m = fmaxf(l0, l1);
      const float e0 = expf(l0 - m), e1 = expf(l1 - m);
      v[i] = e1 / (e0 + e1);
    }
    a.np_map[pix] = lerp(v[0], v[1], v[2], v[3]);
  }
  // horizontal / vertical distance maps, scaled by alpha
  {
    const float* p = a.hv + (size_t)b * 2 * hh;
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      const float* q = p + (size_t)c * hh;
      a.hv_map[pix * 2 + c] = lerp(__ldg(q + o00), __ldg(q + o01), __ldg(q + o10), __ldg(q + o11)) * a.alpha;
    }
  }
  // type probabilities: softmax over K at the four source pixels, interpolate, renormalise
  {
    const float* p = a.tp_logits + (size_t)b * a.K * hh;
    const int o[4] = {o00, o01, o10, o11};
    float mx[4], den[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float m = -INFINITY;
      for (int k = 0; k < a.K; ++k) m = fmaxf(m, __ldg(p + (size_t)k * hh + o[i]));
      float s = 0.f;
      for (int k = 0; k < a.K; ++k) s += expf(__ldg(p + (size_t)k * hh + o[i]) - m);
      mx[i] = m;
      den[i] = s;
    }
    float t[TF_MAXK], tot = 0.f;
    for (int k = 0; k < a.K; ++k) {
      float v[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) v[i] = expf(__ldg(p + (size_t)k * hh + o[i]) - mx[i]) / den[i];
      t[k] = lerp(v[0], v[1], v[2], v[3]);
      tot += t[k];
    }
    const float d = tot + 1e-8f;
    for (int k = 0; k < a.K; ++k) a.tp_map[pix * a.K + k] = t[k] / d;
  }
}

}  // namespace

extern "C" {

int wsi_tilefuse_accumulate(const float* d_np_logits, const float* d_hv, const float* d_tp_logits,
                            const int32_t* d_coords, int coord_stride, int B, int K, int h, int S, int halo, float alpha,
                            int canvas_W, int canvas_H, float* d_np_map, float* d_hv_map, float* d_tp_map,
                            void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  WSI_ARG_CHECK(d_np_logits && d_hv && d_tp_logits && d_coords && d_np_map && d_hv_map && d_tp_map,
                "wsi_tilefuse_accumulate: null pointer");
  WSI_ARG_CHECK(B >= 0 && K >= 1 && K <= TF_MAXK && h >= 1 && S >= 1 && coord_stride >= 2 && canvas_W >= 1 && canvas_H >= 1,
                "wsi_tilefuse_accumulate: bad sizes (K <= %d)", TF_MAXK);
  if (B == 0) return WSI_OK;
  TFArgs a;
  a.np_logits = d_np_logits; a.hv = d_hv; a.tp_logits = d_tp_logits; a.coords = d_coords; a.cstride = coord_stride;
  a.K = K; a.h = h; a.S = S; a.halo = halo; a.alpha = alpha;
  a.scale = (float)h / (float)S;
  a.W = canvas_W; a.H = canvas_H;
  a.np_map = d_np_map; a.hv_map = d_hv_map; a.tp_map = d_tp_map;
  dim3 block(32, 8);
  for (int s = 0; s < B; s += 65535) {
    const int m = B - s < 65535 ? B - s : 65535;
    TFArgs c = a;
    c.B = m;
    c.np_logits += (size_t)s * 2 * h * h;
    c.hv += (size_t)s * 2 * h * h;
    c.tp_logits += (size_t)s * K * h * h;
    c.coords += (size_t)s * coord_stride;
    dim3 grid((S + 31) / 32, (S + 7) / 8, m);
    k_tilefuse<<<grid, block, 0, stream>>>(c);
    wsi_count_launch(1);
  }
  WSI_CUDA_CHECK(cudaGetLastError());
  return WSI_OK;
}

}  // extern "C"
