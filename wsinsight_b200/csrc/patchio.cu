// N4: the optional /images patch cache of the patch file.
//
// Replaces wsinsight/patchlib/io.py:19-48 extract_patches_from_slide (a Python loop of read_region +
// convert("RGB") per patch) with one gather kernel over the device-resident slide, and provides the inverse
// packing (RGB u8 [n][ps][ps][3] -> the RGBA stack the preprocess kernel reads) for
// wsinsight/modellib/data.py:243-268 (_get_patch_from_hdf5).  Pixels outside the slide are zero, as read_region
// pads them (reference tests/test_all.py:747-765).
#include "common.cuh"
#include "../../include/wsi_b200.h"

namespace {

// one thread per output pixel; threadIdx.x runs along the patch row (coalesced 4-byte reads)
__global__ void k_extract_rgb(const uint8_t* __restrict__ slide, long long W, long long H, long long pitch,
                              const int32_t* __restrict__ coords, int ps, uint8_t* __restrict__ out) {
  const int n = blockIdx.z, y = blockIdx.y, x = blockIdx.x * blockDim.x + threadIdx.x;
  if (x >= ps) return;
  const long long sx = (long long)coords[2 * n] + x, sy = (long long)coords[2 * n + 1] + y;
  uint32_t p = 0;
  if (sx >= 0 && sx < W && sy >= 0 && sy < H) p = __ldg(reinterpret_cast<const uint32_t*>(slide + sy * pitch) + sx);
  uint8_t* o = out + (((size_t)n * ps + y) * ps + x) * 3;
  o[0] = (uint8_t)(p & 0xff);
  o[1] = (uint8_t)((p >> 8) & 0xff);
  o[2] = (uint8_t)((p >> 16) & 0xff);
}

// RGB u8 [n][ps][ps][3] -> RGBA u8 [n * ps][pitch] (A = 255, pitch padding zero)
__global__ void k_rgb_to_rgba_stack(const uint8_t* __restrict__ rgb, int ps, long long pitch, uint8_t* __restrict__ out) {
  const int n = blockIdx.z, y = blockIdx.y, x = blockIdx.x * blockDim.x + threadIdx.x;
  if (x >= ps) return;
  const uint8_t* s = rgb + (((size_t)n * ps + y) * ps + x) * 3;
  const uint32_t p = (uint32_t)s[0] | ((uint32_t)s[1] << 8) | ((uint32_t)s[2] << 16) | 0xff000000u;
  reinterpret_cast<uint32_t*>(out + ((size_t)n * ps + y) * pitch)[x] = p;
}

}  // namespace

extern "C" {

int wsi_extract_patches_rgb(const uint8_t* d_slide_rgba, long long W, long long H, long long pitch_bytes,
                            const int32_t* d_coords_xy, int n, int patch_size, uint8_t* d_out_rgb, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  WSI_ARG_CHECK(d_slide_rgba && d_coords_xy && d_out_rgb, "wsi_extract_patches_rgb: null pointer");
  WSI_ARG_CHECK(n >= 0 && patch_size >= 1 && patch_size <= 65535 && (pitch_bytes & 3) == 0 && pitch_bytes >= W * 4,
                "wsi_extract_patches_rgb: bad sizes");
  for (int s = 0; s < n; s += 65535) {        // gridDim.z limit
    const int m = n - s < 65535 ? n - s : 65535;
    dim3 grid((patch_size + 127) / 128, patch_size, m);
    k_extract_rgb<<<grid, 128, 0, stream>>>(d_slide_rgba, W, H, pitch_bytes, d_coords_xy + 2 * (size_t)s, patch_size,
                                            d_out_rgb + (size_t)s * patch_size * patch_size * 3);
    wsi_count_launch(1);
  }
  WSI_CUDA_CHECK(cudaGetLastError());
  return WSI_OK;
}

int wsi_pack_patches_rgba(const uint8_t* d_rgb, int n, int patch_size, long long pitch_bytes, uint8_t* d_out_rgba,
                          void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  WSI_ARG_CHECK(d_rgb && d_out_rgba, "wsi_pack_patches_rgba: null pointer");
  WSI_ARG_CHECK(n >= 0 && patch_size >= 1 && patch_size <= 65535 && (pitch_bytes & 15) == 0 &&
                    pitch_bytes >= (long long)patch_size * 4,
                "wsi_pack_patches_rgba: the stack pitch must be a multiple of 16 bytes and >= 4 * patch_size");
  for (int s = 0; s < n; s += 65535) {
    const int m = n - s < 65535 ? n - s : 65535;
    dim3 grid((patch_size + 127) / 128, patch_size, m);
    k_rgb_to_rgba_stack<<<grid, 128, 0, stream>>>(d_rgb + (size_t)s * patch_size * patch_size * 3, patch_size,
                                                  pitch_bytes, d_out_rgba + (size_t)s * patch_size * pitch_bytes);
    wsi_count_launch(1);
  }
  WSI_CUDA_CHECK(cudaGetLastError());
  return WSI_OK;
}

}  // extern "C"
