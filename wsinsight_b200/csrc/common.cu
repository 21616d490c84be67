// Library-wide helpers: error string, launch counter, TMA descriptor encoding.
#include "common.cuh"
#include "../../include/wsi_b200.h"
#include <atomic>
#include <stdarg.h>

namespace {
thread_local char t_err[1024] = "";
std::atomic<long long> g_launches{0};
}  // namespace

void wsi_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(t_err, sizeof(t_err), fmt, ap);
  va_end(ap);
}

void wsi_count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

wsi_encode_tiled_fn wsi_get_encode_tiled() {
  static wsi_encode_tiled_fn fn = nullptr;
  if (fn) return fn;
  void* p = nullptr;
  cudaDriverEntryPointQueryResult qres;
  cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres);
  if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !p) {
    wsi_set_error("cuTensorMapEncodeTiled not available from the driver (%s)", cudaGetErrorString(e));
    return nullptr;
  }
  fn = (wsi_encode_tiled_fn)p;
  return fn;
}

int wsi_make_tmap(CUtensorMap* out, CUtensorMapDataType dt, int rank, const void* base, const uint64_t* dims,
                  const uint64_t* strides_bytes, const uint32_t* box, const uint32_t* elem_strides,
                  CUtensorMapSwizzle swz) {
  wsi_encode_tiled_fn fn = wsi_get_encode_tiled();
  if (!fn) return WSI_ERR_CUDA;
  cuuint64_t gd[5], gs[4];
  cuuint32_t bx[5], es[5];
  for (int i = 0; i < rank; ++i) {
    gd[i] = dims[i];
    bx[i] = box[i];
    es[i] = elem_strides[i];
  }
  for (int i = 0; i + 1 < rank; ++i) gs[i] = strides_bytes[i];
  CUresult r = fn(out, dt, (cuuint32_t)rank, const_cast<void*>(base), gd, gs, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  swz, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    wsi_set_error(
        "cuTensorMapEncodeTiled failed (CUresult %d): rank %d dims [%llu %llu %llu %llu] strides [%llu %llu %llu] "
        "box [%u %u %u %u] estr [%u %u %u %u] swizzle %d base %p",
        (int)r, rank, (unsigned long long)dims[0], (unsigned long long)(rank > 1 ? dims[1] : 0),
        (unsigned long long)(rank > 2 ? dims[2] : 0), (unsigned long long)(rank > 3 ? dims[3] : 0),
        (unsigned long long)(rank > 1 ? strides_bytes[0] : 0), (unsigned long long)(rank > 2 ? strides_bytes[1] : 0),
        (unsigned long long)(rank > 3 ? strides_bytes[2] : 0), box[0], rank > 1 ? box[1] : 0, rank > 2 ? box[2] : 0,
        rank > 3 ? box[3] : 0, elem_strides[0], rank > 1 ? elem_strides[1] : 0, rank > 2 ? elem_strides[2] : 0,
        rank > 3 ? elem_strides[3] : 0, (int)swz, base);
    return WSI_ERR_CUDA;
  }
  return WSI_OK;
}

extern "C" {

const char* wsi_last_error(void) { return t_err; }

const char* wsi_version(void) { return "wsi_b200 0.1.0 sm_100a"; }

int wsi_memcpy_d2d(void* d_dst, const void* d_src, size_t bytes, void* stream) {
  WSI_ARG_CHECK(d_dst && d_src, "wsi_memcpy_d2d: null pointer");
  WSI_CUDA_CHECK(cudaMemcpyAsync(d_dst, d_src, bytes, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  return WSI_OK;
}

long long wsi_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

// ---- peer-visible result buffer (multi-GPU result path without a gather collective) ----
// Rank 0 owns the (N, K) probability array; the other ranks map it through CUDA IPC and their tail kernel (GAP + FC +
// softmax) stores its rows straight into it over NVLink -- the "compute step followed by a collective" of this path
// becomes the compute step alone.
int wsi_peer_buffer_create(size_t bytes, void** d_ptr, unsigned char* handle64) {
  WSI_ARG_CHECK(d_ptr && handle64 && bytes > 0, "wsi_peer_buffer_create: bad arguments");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
  WSI_CUDA_CHECK(cudaMalloc(d_ptr, bytes));
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, *d_ptr);
  if (e != cudaSuccess) {
    cudaFree(*d_ptr);
    *d_ptr = nullptr;
    wsi_set_error("wsi_peer_buffer_create: cudaIpcGetMemHandle -> %s", cudaGetErrorString(e));
    return WSI_ERR_CUDA;
  }
  memcpy(handle64, &h, 64);
  return WSI_OK;
}

int wsi_peer_buffer_open(const unsigned char* handle64, void** d_ptr) {
  WSI_ARG_CHECK(d_ptr && handle64, "wsi_peer_buffer_open: bad arguments");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  WSI_CUDA_CHECK(cudaIpcOpenMemHandle(d_ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return WSI_OK;
}

int wsi_peer_buffer_close(void* d_ptr) {
  WSI_CUDA_CHECK(cudaIpcCloseMemHandle(d_ptr));
  return WSI_OK;
}

int wsi_peer_buffer_free(void* d_ptr) {
  WSI_CUDA_CHECK(cudaFree(d_ptr));
  return WSI_OK;
}

}  // extern "C"
