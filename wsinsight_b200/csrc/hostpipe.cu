// Host-buffer entry of the classification loop: pinned RGBA patches in, fp32 probabilities out.
//
// Replaces one iteration of the reference's classification loop
// (wsinsight/modellib/run_inference.py:539-556): the DataLoader hands over a HOST batch, the
// reference does `model(batch.to(device)).detach().cpu()` and a CPU softmax.  Here the host
// batch is the raw RGBA pixels `read_region` returned (wsinsight/modellib/data.py:276-280),
// stacked patch after patch; the copy to the device, the fused crop/resize/normalise kernel,
// the network forward (softmax in the tail kernel) and the copy back of the probabilities
// are enqueued on two streams over `depth` staging slots, so the H2D copy of batch i+1
// overlaps the kernels of batch i.  Nothing here computes on the CPU.
#include "common.cuh"
#include "../../include/wsi_b200.h"
#include <stdlib.h>
#include <vector>

namespace {

constexpr int HP_CHUNKS = 4;    // a batch is copied and preprocessed in this many pieces (pipeline fill = one piece)

struct Slot {
  uint8_t* d_rgb = nullptr;            // RGB host mode: [max_batch][ps][ps][3] u8 landing buffer (+ 16 bytes of slack)
  uint8_t* d_stage = nullptr;          // [max_batch * ps rows][pitch] RGBA u8
  __nv_bfloat16* d_input = nullptr;    // network input layout, borders zeroed once
  float* d_probs = nullptr;            // [max_batch][K]
  void* plan = nullptr;                // preprocess plan over d_stage seen as a (ps x max_batch*ps) image
  cudaEvent_t copied = nullptr;        // H2D finished
  cudaEvent_t chunk_ev[HP_CHUNKS] = {nullptr, nullptr, nullptr, nullptr};   // H2D of piece c finished
  cudaEvent_t staged_free = nullptr;   // preprocess finished reading d_stage
  cudaEvent_t done = nullptr;          // D2H of the probabilities finished
  bool used = false;
};

struct HostPipe {
  void* net = nullptr;
  int ps = 0, S = 0, K = 0, max_batch = 0, depth = 0;
  long long pitch = 0;
  int in_H = 0, in_W = 0;
  bool skip_h2d = false;
  int host_channels = 4;               // 4: RGBA rows of `pitch` bytes (what read_region returns); 3: dense RGB
  long long patch_elems = 0;           // bf16 elements one patch occupies in the network input buffer (per plane)
  int32_t* d_coords = nullptr;         // (0, i*ps) for i < max_batch
  cudaStream_t s_copy = nullptr, s_comp = nullptr;
  std::vector<Slot> slots;
  long long submitted = 0;
  size_t h2d_bytes = 0, d2h_bytes = 0;
};

void destroy(HostPipe* p) {
  if (!p) return;
  for (auto& s : p->slots) {
    if (s.plan) wsi_preproc_plan_destroy(s.plan);
    if (s.d_stage) cudaFree(s.d_stage);
    if (s.d_rgb) cudaFree(s.d_rgb);
    if (s.d_input) cudaFree(s.d_input);
    if (s.d_probs) cudaFree(s.d_probs);
    if (s.copied) cudaEventDestroy(s.copied);
    for (auto e : s.chunk_ev)
      if (e) cudaEventDestroy(e);
    if (s.staged_free) cudaEventDestroy(s.staged_free);
    if (s.done) cudaEventDestroy(s.done);
  }
  if (p->d_coords) cudaFree(p->d_coords);
  if (p->s_copy) cudaStreamDestroy(p->s_copy);
  if (p->s_comp) cudaStreamDestroy(p->s_comp);
  delete p;
}

}  // namespace

extern "C" {

int wsi_hostpipe_create(void* net, int patch_size, long long patch_pitch_bytes, int out_size, const float* mean,
                        const float* stdv, int max_batch, int depth, void** pipe_out) {
  WSI_ARG_CHECK(net && pipe_out, "wsi_hostpipe_create: null pointer");
  WSI_ARG_CHECK(patch_size >= 1 && max_batch >= 1 && depth >= 1 && depth <= 8, "wsi_hostpipe_create: bad sizes");
  WSI_ARG_CHECK((patch_pitch_bytes & 15) == 0 && patch_pitch_bytes >= (long long)patch_size * 4,
                "wsi_hostpipe_create: the row pitch of a host patch must be a multiple of 16 bytes and >= 4*patch_size");
  int hwc[3] = {0, 0, 0};
  int rc = wsi_net_input_geometry(net, hwc);
  if (rc) return rc;
  WSI_ARG_CHECK((hwc[2] == 4 || hwc[2] == 8) && hwc[0] >= out_size && hwc[1] >= out_size && hwc[0] == hwc[1],
                "wsi_hostpipe_create: network input %dx%dx%d does not fit out_size %d", hwc[0], hwc[1], hwc[2],
                out_size);
  HostPipe* p = new HostPipe();
  p->net = net;
  p->ps = patch_size;
  p->S = out_size;
  p->K = wsi_net_num_classes(net);
  p->max_batch = max_batch;
  p->depth = depth;
  p->pitch = patch_pitch_bytes;
  {
    const char* e = getenv("WSI_B200_HP_NO_H2D");
    p->skip_h2d = e && e[0] == '1';
  }
  p->in_H = hwc[0];
  p->in_W = hwc[1];
  // symmetric stem padding (ResNet: 3); an odd difference is the one unused row/column that rounds an odd
  // input size up to the even side the parity-plane layout needs (InceptionV4: 299 -> 300, pad 0)
  const int pad = (hwc[0] - out_size) / 2;
  auto fail = [&](const char* what, int code) {
    wsi_set_error("wsi_hostpipe_create: %s", what);
    destroy(p);
    return code;
  };
  if (cudaStreamCreateWithFlags(&p->s_copy, cudaStreamNonBlocking) != cudaSuccess ||
      cudaStreamCreateWithFlags(&p->s_comp, cudaStreamNonBlocking) != cudaSuccess)
    return fail("stream creation failed", WSI_ERR_CUDA);
  std::vector<int32_t> hc((size_t)max_batch * 2);
  for (int i = 0; i < max_batch; ++i) {
    hc[2 * i] = 0;
    hc[2 * i + 1] = i * patch_size;
  }
  if (cudaMalloc(&p->d_coords, hc.size() * sizeof(int32_t)) != cudaSuccess) return fail("cudaMalloc", WSI_ERR_NOMEM);
  if (cudaMemcpy(p->d_coords, hc.data(), hc.size() * sizeof(int32_t), cudaMemcpyHostToDevice) != cudaSuccess)
    return fail("copy of the patch-slot coordinates failed", WSI_ERR_CUDA);
  p->slots.resize(depth);
  const size_t stage_bytes = (size_t)max_batch * patch_size * (size_t)patch_pitch_bytes;
  const size_t input_bytes = wsi_net_input_bytes(net);
  int lay[4] = {0, 0, 0, 0};
  rc = wsi_net_input_layout(net, lay);
  if (rc) {
    destroy(p);
    return rc;
  }
  const long long plane_stride_px = lay[0] == 1 ? (((long long)lay[3] << 32) | (unsigned int)lay[2]) : 0;
  p->patch_elems = plane_stride_px ? (long long)(hwc[0] / 2) * hwc[1] * 4 : (long long)hwc[0] * hwc[1] * hwc[2];
  for (auto& s : p->slots) {
    if (cudaMalloc(&s.d_stage, stage_bytes) != cudaSuccess || cudaMalloc(&s.d_input, input_bytes) != cudaSuccess ||
        cudaMalloc(&s.d_probs, (size_t)max_batch * p->K * sizeof(float)) != cudaSuccess)
      return fail("cudaMalloc", WSI_ERR_NOMEM);
    if (cudaMemset(s.d_input, 0, input_bytes) != cudaSuccess) return fail("cudaMemset of the input buffer failed", WSI_ERR_CUDA);
    if (cudaEventCreateWithFlags(&s.copied, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&s.staged_free, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&s.done, cudaEventDisableTiming) != cudaSuccess)
      return fail("event creation failed", WSI_ERR_CUDA);
    for (auto& e : s.chunk_ev)
      if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess)
        return fail("event creation failed", WSI_ERR_CUDA);
    rc = wsi_preproc_plan_create(s.d_stage, patch_size, (long long)max_batch * patch_size, patch_pitch_bytes,
                                 patch_size, out_size, mean, stdv, hwc[0], hwc[1], pad, pad, &s.plan);
    if (!rc) rc = wsi_preproc_plan_set_parity_layout(s.plan, plane_stride_px);
    if (!rc) rc = wsi_preproc_plan_set_channels(s.plan, hwc[2]);
    if (rc) {
      destroy(p);
      return rc;
    }
  }
  if (cudaDeviceSynchronize() != cudaSuccess) return fail("cudaDeviceSynchronize failed", WSI_ERR_CUDA);
  *pipe_out = p;
  return WSI_OK;
}

int wsi_hostpipe_set_host_channels(void* pipe, int channels) {
  HostPipe* p = (HostPipe*)pipe;
  WSI_ARG_CHECK(p && (channels == 3 || channels == 4), "wsi_hostpipe_set_host_channels: channels must be 3 or 4");
  WSI_ARG_CHECK(p->submitted == 0, "wsi_hostpipe_set_host_channels: call before the first submit");
  if (channels == 3)
    for (auto& s : p->slots)
      if (!s.d_rgb) {
        const size_t bytes = (size_t)p->max_batch * p->ps * p->ps * 3 + 64;
        if (cudaMalloc(&s.d_rgb, bytes) != cudaSuccess) {
          wsi_set_error("wsi_hostpipe_set_host_channels: cudaMalloc of the RGB landing buffer failed");
          return WSI_ERR_NOMEM;
        }
        // the RGBA stack's pitch padding (columns >= ps) is never written by the pack kernel: clear it once
        WSI_CUDA_CHECK(cudaMemset(s.d_stage, 0, (size_t)p->max_batch * p->ps * (size_t)p->pitch));
      }
  p->host_channels = channels;
  return WSI_OK;
}

int wsi_hostpipe_submit(void* pipe, const uint8_t* h_rgba, int n, float* h_probs) {
  HostPipe* p = (HostPipe*)pipe;
  WSI_ARG_CHECK(p && h_rgba && h_probs, "wsi_hostpipe_submit: null pointer");
  WSI_ARG_CHECK(n >= 1 && n <= p->max_batch, "wsi_hostpipe_submit: batch %d out of range (max %d)", n, p->max_batch);
  Slot& s = p->slots[(size_t)(p->submitted % p->depth)];
  const size_t in_bytes = p->host_channels == 3 ? (size_t)n * p->ps * p->ps * 3 : (size_t)n * p->ps * (size_t)p->pitch;
  const size_t out_bytes = (size_t)n * p->K * sizeof(float);
  if (s.used) {
    // the staging buffer may be overwritten once the preprocess kernel that read it has finished;
    // the probabilities of that earlier batch must have left d_probs before it is rewritten
    WSI_CUDA_CHECK(cudaStreamWaitEvent(p->s_copy, s.staged_free, 0));
  }
  // The batch goes over in HP_CHUNKS pieces, each preprocessed as soon as it has landed: the copy of piece c + 1 overlaps
  // the preprocessing of piece c, and what a cold pipeline exposes is one piece, not one batch (at N = 8 the first
  // 126-MB copies of all ranks contend for the host's memory and PCIe root and were 5 % of a 20-step window).
  const int C = n >= 64 * HP_CHUNKS ? HP_CHUNKS : 1;
  const int per = (n + C - 1) / C;
  const size_t patch_bytes = (size_t)p->ps * (size_t)p->pitch;
  const size_t rgb_bytes = (size_t)p->ps * p->ps * 3;
  int rc = WSI_OK;
  for (int c = 0, j0 = 0; j0 < n; ++c, j0 += per) {
    const int m = n - j0 < per ? n - j0 : per;
    if (p->host_channels == 3) {
      // dense RGB patches (25 % fewer bytes over PCIe): land in d_rgb, expanded to the RGBA stack on the device
      if (!p->skip_h2d)
        WSI_CUDA_CHECK(cudaMemcpyAsync(s.d_rgb + (size_t)j0 * rgb_bytes, h_rgba + (size_t)j0 * rgb_bytes,
                                       (size_t)m * rgb_bytes, cudaMemcpyHostToDevice, p->s_copy));
    } else if (!p->skip_h2d) {   // WSI_B200_HP_NO_H2D=1: timing experiment only (stale pixels), never set in production
      WSI_CUDA_CHECK(cudaMemcpyAsync(s.d_stage + (size_t)j0 * patch_bytes, h_rgba + (size_t)j0 * patch_bytes,
                                     (size_t)m * patch_bytes, cudaMemcpyHostToDevice, p->s_copy));
    }
    if (p->host_channels == 3) {
      // expansion to the RGBA stack on the COPY stream, right behind the piece's H2D: it runs in the shadow of the
      // previous batch's convolutions instead of sitting on the compute stream's critical path
      rc = wsi_pack_patches_rgba_fast(s.d_rgb, j0, m, p->ps, p->pitch, s.d_stage, p->s_copy);
      if (rc) return rc;
    }
    WSI_CUDA_CHECK(cudaEventRecord(s.chunk_ev[c], p->s_copy));
    WSI_CUDA_CHECK(cudaStreamWaitEvent(p->s_comp, s.chunk_ev[c], 0));
    rc = wsi_preproc_run(s.plan, p->d_coords + 2 * (size_t)j0, m, s.d_input + (size_t)j0 * p->patch_elems, p->s_comp);
    if (rc) return rc;
  }
  WSI_CUDA_CHECK(cudaEventRecord(s.copied, p->s_copy));
  WSI_CUDA_CHECK(cudaEventRecord(s.staged_free, p->s_comp));
  rc = wsi_net_forward(p->net, s.d_input, n, s.d_probs, p->s_comp);
  if (rc) return rc;
  WSI_CUDA_CHECK(cudaMemcpyAsync(h_probs, s.d_probs, out_bytes, cudaMemcpyDeviceToHost, p->s_comp));
  WSI_CUDA_CHECK(cudaEventRecord(s.done, p->s_comp));
  s.used = true;
  p->submitted++;
  p->h2d_bytes += in_bytes;
  p->d2h_bytes += out_bytes;
  return WSI_OK;
}

int wsi_hostpipe_sync(void* pipe) {
  HostPipe* p = (HostPipe*)pipe;
  WSI_ARG_CHECK(p, "wsi_hostpipe_sync: null pointer");
  WSI_CUDA_CHECK(cudaStreamSynchronize(p->s_copy));
  WSI_CUDA_CHECK(cudaStreamSynchronize(p->s_comp));
  return WSI_OK;
}

void* wsi_hostpipe_compute_stream(void* pipe) {
  HostPipe* p = (HostPipe*)pipe;
  return p ? (void*)p->s_comp : nullptr;
}

int wsi_hostpipe_bytes(void* pipe, long long* h2d_bytes, long long* d2h_bytes) {
  HostPipe* p = (HostPipe*)pipe;
  WSI_ARG_CHECK(p, "wsi_hostpipe_bytes: null pointer");
  if (h2d_bytes) *h2d_bytes = (long long)p->h2d_bytes;
  if (d2h_bytes) *d2h_bytes = (long long)p->d2h_bytes;
  return WSI_OK;
}

void wsi_hostpipe_destroy(void* pipe) { destroy((HostPipe*)pipe); }

int wsi_host_alloc_pinned(size_t bytes, void** h_ptr) {
  WSI_ARG_CHECK(h_ptr, "wsi_host_alloc_pinned: null pointer");
  WSI_CUDA_CHECK(cudaHostAlloc(h_ptr, bytes, cudaHostAllocDefault));
  return WSI_OK;
}

int wsi_host_free_pinned(void* h_ptr) {
  WSI_CUDA_CHECK(cudaFreeHost(h_ptr));
  return WSI_OK;
}

}  // extern "C"
