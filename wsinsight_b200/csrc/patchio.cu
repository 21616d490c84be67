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

// Same packing at memory speed (the host-buffer pipeline runs it on every batch): a CTA takes PK_ROWS rows of one
// patch -- a contiguous run of bytes in the dense RGB stack -- loads it with aligned 16-byte vectors into shared
// memory, and writes one RGBA word per pixel, coalesced along the row.
constexpr int PK_ROWS = 16;

__global__ void __launch_bounds__(256)
    k_rgb_to_rgba_rows(const uint8_t* __restrict__ rgb, int patch0, int ps, long long pitch, uint8_t* __restrict__ out) {
  extern __shared__ __align__(16) uint8_t s_pk[];
  const int n = patch0 + blockIdx.y, y0 = blockIdx.x * PK_ROWS;
  const int rows = min(PK_ROWS, ps - y0);
  const size_t first = ((size_t)n * ps + y0) * (size_t)ps * 3;       // first byte of the run
  const size_t nbytes = (size_t)rows * ps * 3;
  const size_t a0 = first & ~(size_t)15;                               // aligned load window
  const int mis = (int)(first - a0);
  const int nvec = (int)((mis + nbytes + 15) / 16);
  const uint4* src = reinterpret_cast<const uint4*>(rgb + a0);
  uint4* dst = reinterpret_cast<uint4*>(s_pk);
  // the window may start up to 15 bytes before / end up to 15 bytes after the run: still inside the allocation except
  // at the very end of the stack, where the caller guarantees 16 bytes of slack
  for (int i = threadIdx.x; i < nvec; i += 256) dst[i] = __ldg(src + i);
  __syncthreads();
  const int npx = rows * ps;
  for (int p = threadIdx.x; p < npx; p += 256) {
    const int y = p / ps, x = p - y * ps;
    const uint8_t* q = s_pk + mis + 3 * p;
    const uint32_t v = (uint32_t)q[0] | ((uint32_t)q[1] << 8) | ((uint32_t)q[2] << 16) | 0xff000000u;
    reinterpret_cast<uint32_t*>(out + ((size_t)n * ps + y0 + y) * pitch)[x] = v;
  }
}

// Macenko stain normalisation of cached patches (N4): HistomicsTK's deconvolution_based_normalization(patch + 1e-8,
// W_source, W_target) per pixel (wsinsight/modellib/data.py:293-300).  With x = (float)(u8 + 1e-8f) + 1 (the library's
// optical-density form: white = 256), sda = -log(x / 256) * 255 / log(256); stains = inv(Wc) . sda; the library then
// maps the stains back to "RGB" (256^(1 - s/255) - 1), forward again (+1, log) and multiplies by W_target: the round
// trip is the identity up to rounding, so M = W_target . inv(Wc) is applied once.  out = trunc(clip(256^(1 - v/255) - 1,
// 0, 255)).  float64 throughout, as numpy computes it.
struct StainM { double m[9]; };

__global__ void __launch_bounds__(256)
    k_stain_normalize(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, long long npix, StainM M) {
  const long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= npix) return;
  const double k = 255.0 / log(256.0);
  double sda[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const float f = (float)in[3 * p + c] + 1e-8f;          // patch_arr.astype(float32) + EPSILON
    const double x = fmax((double)f + 1.0, 1e-10);
    sda[c] = -log(x / 256.0) * k;
  }
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const double v = M.m[3 * c] * sda[0] + M.m[3 * c + 1] * sda[1] + M.m[3 * c + 2] * sda[2];
    double r = pow(256.0, 1.0 - v / 255.0) - 1.0;
    r = fmin(fmax(r, 0.0), 255.0);
    out[3 * p + c] = (uint8_t)r;                             // astype(np.uint8): truncation
  }
}

}  // namespace

extern "C" {

// Patches [first_patch, first_patch + n) of the dense RGB stack d_rgb ([..][ps][ps][3], base 16-byte aligned and followed
// by >= 16 readable bytes: the aligned vector loads of the last run) -> the same patch slots of the RGBA stack.
int wsi_pack_patches_rgba_fast(const uint8_t* d_rgb, int first_patch, int n, int patch_size, long long pitch_bytes,
                               uint8_t* d_out_rgba, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  WSI_ARG_CHECK(d_rgb && d_out_rgba, "wsi_pack_patches_rgba_fast: null pointer");
  WSI_ARG_CHECK(first_patch >= 0 && n >= 0 && n <= 65535 && patch_size >= 1 && patch_size <= 1024 && (pitch_bytes & 15) == 0 &&
                    pitch_bytes >= (long long)patch_size * 4 && (((uintptr_t)d_rgb) & 15) == 0,
                "wsi_pack_patches_rgba_fast: bad sizes / alignment");
  if (n == 0) return WSI_OK;
  const size_t smem = (size_t)PK_ROWS * patch_size * 3 + 32;
  static bool attr_set = false;
  if (!attr_set) {
    WSI_CUDA_CHECK(cudaFuncSetAttribute(k_rgb_to_rgba_rows, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
    attr_set = true;
  }
  dim3 grid((patch_size + PK_ROWS - 1) / PK_ROWS, n);
  k_rgb_to_rgba_rows<<<grid, 256, smem, stream>>>(d_rgb, first_patch, patch_size, pitch_bytes, d_out_rgba);
  wsi_count_launch(1);
  WSI_CUDA_CHECK(cudaGetLastError());
  return WSI_OK;
}

int wsi_stain_normalize_rgb(const uint8_t* d_rgb_in, uint8_t* d_rgb_out, long long n_pixels, const double* h_M9,
                            void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  WSI_ARG_CHECK(d_rgb_in && d_rgb_out && h_M9 && n_pixels >= 0, "wsi_stain_normalize_rgb: bad arguments");
  if (n_pixels == 0) return WSI_OK;
  StainM M;
  for (int i = 0; i < 9; ++i) M.m[i] = h_M9[i];
  const long long nb = (n_pixels + 255) / 256;
  WSI_ARG_CHECK(nb < (1ll << 31), "wsi_stain_normalize_rgb: too many pixels for one call");
  k_stain_normalize<<<(unsigned)nb, 256, 0, stream>>>(d_rgb_in, d_rgb_out, n_pixels, M);
  wsi_count_launch(1);
  WSI_CUDA_CHECK(cudaGetLastError());
  return WSI_OK;
}

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
