/* wsi_b200.h -- C ABI of libwsi_b200.so: the sm_100a kernels behind the WSInsight
 * patch-classification hot path (tissue segmentation + patchify -> crop/resize/normalize ->
 * CNN forward -> per-patch probabilities).
 *
 * The reference (huangch/wsinsight) is pure Python and has no FFI of its own; each entry
 * point below names the reference call site (file:line under /root/reference) whose
 * arithmetic it replaces.  The Python host layer (wsinsight_b200/patchlib, modellib) binds
 * these with ctypes and mirrors the reference's Python signatures; INTEGRATION.md shows the
 * stub a maintainer of the reference would add.
 *
 * Conventions
 *   - plain pointers and sizes only; every `d_` pointer is DEVICE memory, every `h_`
 *     pointer is host memory; `stream` is a cudaStream_t passed as void*.
 *   - return 0 on success, a negative WSI_ERR_* code otherwise; the message is
 *     retrievable with wsi_last_error() (thread-local).  Nothing throws.
 *   - work is enqueued on `stream` and NOT synchronised unless stated.
 *   - there is no CPU fallback anywhere in this library.
 */
#ifndef WSI_B200_H_
#define WSI_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define WSI_OK 0
#define WSI_ERR_CUDA (-1)
#define WSI_ERR_ARG (-2)
#define WSI_ERR_UNSUPPORTED (-3)
#define WSI_ERR_NOMEM (-4)

/* ---- library ------------------------------------------------------------------------- */
const char* wsi_last_error(void);
/* "wsi_b200 <version> sm_100a" */
const char* wsi_version(void);
/* number of kernels this library has launched in this process (all streams) */
long long wsi_launch_count(void);

/* device-to-device copy on `stream` (debug access to activations owned by the library) */
int wsi_memcpy_d2d(void* d_dst, const void* d_src, size_t bytes, void* stream);

/* Peer-visible result buffer: the multi-GPU result path without a gather collective.  The reference gathers per-GPU
 * outputs inside nn.DataParallel (wsinsight/modellib/run_inference.py:173-174).  Here rank 0 creates the (N, K) fp32
 * probability array, the other ranks of the node open it through CUDA IPC (handle64: 64 opaque bytes, sent by any
 * side channel) and pass `mapped + row offset` as the output pointer of wsi_net_forward: the tail kernel stores its
 * probabilities straight into rank 0's memory over NVLink. */
int wsi_peer_buffer_create(size_t bytes, void** d_ptr, unsigned char* handle64);
int wsi_peer_buffer_open(const unsigned char* handle64, void** d_ptr);
int wsi_peer_buffer_close(void* d_ptr);
int wsi_peer_buffer_free(void* d_ptr);

/* ---- K1: tissue segmentation --------------------------------------------------------- */
/* Replaces segment_tissue(), wsinsight/patchlib/segment.py:13-97:
 *   cv.cvtColor(RGB2HSV)[...,1] (:54,65) -> cv.medianBlur(k) (:77) -> cv.threshold (:80-83)
 *   -> skimage binary_closing(ones(c,c)) (:87-89) -> remove_small_objects (:92)
 *   -> remove_small_holes (:95).
 * d_rgb: H x W x 3 u8, row pitch `pitch_bytes`.  d_mask: H x W u8 (0/1), the function's result.
 * d_outer (optional, may be NULL): H x W u8, mask restricted to the OUTERMOST 8-connected
 * components (what get_multipolygon_from_binary_arr, patchlib/patch.py:35-130, keeps after its
 * sequential union/difference of cv.findContours(RETR_CCOMP) contours); input of wsi_patch_grid.
 * Errors: median_filter_size <= 1 or even -> WSI_ERR_ARG with the reference's ValueError text
 * (segment.py:68-72). */
size_t wsi_seg_workspace_bytes(int H, int W);
int wsi_seg_tissue_u8(const uint8_t* d_rgb, int H, int W, int pitch_bytes, int median_filter_size,
                      int binary_threshold, int closing_kernel_size, int min_object_size_px,
                      int min_hole_size_px, uint8_t* d_mask, uint8_t* d_outer, void* d_ws,
                      size_t ws_bytes, void* stream);

/* Batched form for a cohort of thumbnails (BASELINE config 4): slide s has geometry
 * (h_H[s], h_W[s], h_pitch[s]) and lives at d_rgb + h_rgb_off[s]; mask/outer at
 * d_mask + h_mask_off[s].  Thresholds are per slide (they depend on the thumbnail mpp,
 * wsinsight/patchlib/pipeline.py:107-112).  Every stage is ONE launch over all slides. */
size_t wsi_seg_batch_workspace_bytes(int n, const int* h_H, const int* h_W);
int wsi_seg_tissue_batch_u8(const uint8_t* d_rgb, int n, const int* h_H, const int* h_W,
                            const int* h_pitch, const long long* h_rgb_off,
                            const long long* h_mask_off, int median_filter_size,
                            int binary_threshold, int closing_kernel_size,
                            const int* h_min_object_px, const int* h_min_hole_px, uint8_t* d_mask,
                            uint8_t* d_outer, void* d_ws, size_t ws_bytes, void* stream);

/* Extended form.  Two things the reference leaves to its dependencies are parameters here:
 *  - binary_threshold == WSI_SEG_THRESHOLD_OTSU: the threshold of cv.threshold (segment.py:80-83) is chosen per
 *    slide by Otsu's method on the median-filtered saturation image, bit-identical to
 *    cv.threshold(S_med, 0, 255, THRESH_BINARY | THRESH_OTSU) (OpenCV getThreshVal_Otsu_8u).  NOT upstream
 *    behaviour (the reference always uses the fixed threshold, default 7): an extra mode named by the task.
 *  - closing_dilation_origin / closing_erosion_origin: the column/row of the k x k footprint that sits on the
 *    output pixel (window offsets [-origin .. k-1-origin]) for the dilation and the erosion of
 *    binary_closing(ones(k,k)) (segment.py:87-89).  -1 selects the default, the raw even footprint as
 *    scipy.ndimage centres it: dilation (k-1)/2, erosion k/2 (k = 6: offsets -2..+3 then -3..+2).  A footprint
 *    front-padded to (k+1) x (k+1) is dilation k/2, erosion (k-1)/2.  The conventions agree for odd k and,
 *    for even k, everywhere except within k/2 pixels of the image border.
 * d_info (optional): int32 [n][4] per slide {threshold used, mask has tissue (valid when d_outer != NULL), 0, 0}. */
#define WSI_SEG_THRESHOLD_OTSU (-1)
int wsi_seg_tissue_batch_ex(const uint8_t* d_rgb, int n, const int* h_H, const int* h_W,
                            const int* h_pitch, const long long* h_rgb_off,
                            const long long* h_mask_off, int median_filter_size,
                            int binary_threshold, int closing_kernel_size,
                            int closing_dilation_origin, int closing_erosion_origin,
                            const int* h_min_object_px, const int* h_min_hole_px, uint8_t* d_mask,
                            uint8_t* d_outer, int* d_info, void* d_ws, size_t ws_bytes, void* stream);

/* Outermost-component restriction of an existing 0/1 mask (the polygon stage alone):
 * get_multipolygon_from_binary_arr(), wsinsight/patchlib/patch.py:35-130. */
int wsi_outermost_u8(const uint8_t* d_mask, int H, int W, uint8_t* d_outer, void* d_ws,
                     size_t ws_bytes, void* stream);

/* ---- K2: in-tissue patch grid --------------------------------------------------------- */
/* Replaces get_patch_coordinates_within_polygon(), wsinsight/patchlib/patch.py:174-242
 * (itertools grid :218-225, STRtree.query(contains) :227-233, ordering :236, -half :240) and the
 * polygon it is given (patch.py:57-130).  Candidates are centroids (half + i*step, half + j*step)
 * with half + i*step < slide_w, half + j*step < slide_h; kept candidates are written as top-left
 * (x, y) int32 pairs in the reference's order (x-major, y fastest).  *d_count receives the number
 * kept (may exceed `capacity`; only `capacity` pairs are written).  h_tile_dim[2] (optional)
 * receives (n_i, n_j) == tile_dim of wsinsight/patchlib/pipeline.py:370-382. */
size_t wsi_patch_grid_workspace_bytes(long long slide_w, long long slide_h, int half, int step);
int wsi_patch_grid(const uint8_t* d_outer, int Ht, int Wt, long long slide_w, long long slide_h,
                   int half, int step, int32_t* d_coords_xy, int capacity, int* d_count,
                   int* h_tile_dim, void* d_ws, size_t ws_bytes, void* stream);
/* Allocation-free variant of wsi_patch_grid.  NO counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  The scratch lives in a file-scope `static __device__` array of the library (256 KB per GPU that loads the
 * library, held until exit) sized for at most WSI_PATCH_GRID_STATIC_MAX_CANDIDATES candidate centroids
 * (ni * nj of wsi_patch_grid); larger grids return WSI_ERR_ARG.  Inputs, outputs and results are those of
 * wsi_patch_grid.  The call performs no cudaMalloc / cudaFree / allocating library call.  All calls on one device share
 * the array: calls that overlap on the GPU (different streams) race -- keep them on one stream or serialise them.
 * The first call on each device resolves the array's device address; make that call outside a graph capture. */
#define WSI_PATCH_GRID_STATIC_MAX_CANDIDATES (1 << 24)
int wsi_patch_grid_static(const uint8_t* d_outer, int Ht, int Wt, long long slide_w, long long slide_h,
                          int half, int step, int32_t* d_coords_xy, int capacity, int* d_count,
                          int* h_tile_dim, void* stream);

/* ---- K3: fused patch data path --------------------------------------------------------- */
/* Replaces WholeSlideImagePatches._get_patch_from_wsi (wsinsight/modellib/data.py:270-281:
 * read_region -> RGBA, zero padded out of bounds -> convert("RGB")), the torchvision Compose of
 * wsinsight/modellib/transforms.py:22-38 (Resize = PIL antialiased bilinear, ToTensor, Normalize)
 * and the default collate + H2D copy of wsinsight/modellib/run_inference.py:288-299,543.
 * The slide is a device-resident RGBA u8 image (16-byte aligned base and pitch).  Output:
 * bf16 [n][out_H][out_W][4] (channel 3 = 0) with the S x S patch at (pad_top, pad_left); border
 * elements are never written (zero the buffer once).  mean/stdv may both be NULL (no Normalize). */
int wsi_preproc_plan_create(const uint8_t* d_slide_rgba, long long slide_w, long long slide_h,
                            long long pitch_bytes, int patch_size, int out_size,
                            const float* mean, const float* stdv, int out_H, int out_W,
                            int pad_top, int pad_left, void** plan_out);
int wsi_preproc_run(void* plan, const int32_t* d_coords_xy, int n, void* d_out_bf16, void* stream);
/* debug/parity: the u8 result of the resize alone, [n][S][S][4] (RGB0) */
int wsi_preproc_run_u8(void* plan, const int32_t* d_coords_xy, int n, uint8_t* d_out_u8,
                       void* stream);
/* plane_stride_px > 0 switches the output to the layout the parity-plane stem kernel reads (see
 * wsi_net_input_layout): two row-parity planes, bf16 [2][max_batch][P][2P][4] with P = out_H/2, planes
 * plane_stride_px pixels apart; 0 restores [n][out_H][out_W][4]. */
int wsi_preproc_plan_set_parity_layout(void* plan, long long plane_stride_px);
/* Dense layout only: the number of patches the output buffer passed to wsi_preproc_run holds (0 = unchecked).
 * wsi_preproc_run returns WSI_ERR_ARG when n exceeds it; with the parity layout the capacity follows from
 * plane_stride_px and is always checked (a larger batch would spill plane 0 into plane 1). */
int wsi_preproc_plan_set_capacity(void* plan, int max_patches);
/* bf16 channels per output pixel: 4 (default) or 8 (first convolutions that run on the generic kernel);
 * channels 3.. are zero. */
int wsi_preproc_plan_set_channels(void* plan, int channels);
int wsi_preproc_plan_info(void* plan, int* info8);
void wsi_preproc_plan_destroy(void* plan);

/* ---- K4: CNN forward ------------------------------------------------------------------- */
/* Replaces `model(batch_imgs.to(device))` + CPU softmax/sigmoid of the classification loop,
 * wsinsight/modellib/run_inference.py:537-560, for zoo CNNs whose TorchScript graph
 * (wsinsight/modellib/models.py:51-108) has been lowered by the host to a list of
 * conv(+bias,+residual,+ReLU) / pool / head ops over NHWC bf16 tensors.  Convolutions are
 * tcgen05/TMEM implicit GEMMs fed by TMA; weights are given as host fp32 OIHW with BatchNorm
 * already folded (what torch.jit.freeze leaves, scripts/convert_to_torchscript.py:38-50). */
int wsi_net_create(int max_batch, void** net_out);
int wsi_net_add_tensor(void* net, int H, int W, int C, int external, int* id_out);
int wsi_net_add_conv(void* net, int in_id, int out_id, int out_c_off, const float* h_w_oihw,
                     const float* h_bias, int Cin, int Cout, int KH, int KW, int stride, int pad_h,
                     int pad_w, int relu, int residual_id);
/* Same, reading input channels [in_c_off, in_c_off + Cin) of a wider internal tensor: the engine merges sibling 1x1
 * convolutions of an Inception block (same input: `torch.cat` branches of the TorchScript graph the reference runs,
 * run_inference.py:541) into one launch writing one tensor, and the branch convolutions that follow read slices. */
int wsi_net_add_conv_slice(void* net, int in_id, int in_c_off, int out_id, int out_c_off, const float* h_w_oihw,
                           const float* h_bias, int Cin, int Cout, int KH, int KW, int stride, int pad_h, int pad_w,
                           int relu, int residual_id);
int wsi_net_add_stem7x7(void* net, int in_id, int out_id, const float* h_w_oihw,
                        const float* h_bias, int Cout, int relu);
/* Generalisation: ksize 7 / pad 3 (ResNet) or ksize 3 / pad 0 (InceptionV4 conv2d_1a; input tensor [S'][S'][4] with
 * S' = S rounded up to even), stride 2.  Replaces the first cuDNN convolution of the TorchScript graph
 * (modellib/run_inference.py:543). */
int wsi_net_add_stem_parity(void* net, int in_id, int out_id, const float* w_oihw, const float* bias, int Cout,
                            int ksize, int pad, int relu);
int wsi_net_add_pool(void* net, int in_id, int out_id, int out_c_off, int is_max, int k, int stride,
                     int pad, int count_include_pad);
/* 3x3 / stride-1 average pool followed by + bias[c] and optional ReLU: the pool branch of an Inception block
 * (avg_pool2d -> 1x1 BasicConv2d in the TorchScript graph, executed by run_inference.py:541) is run as
 * 1x1 conv (no bias, no ReLU) -> this op, which is the same linear map with the pooling pass over 4-8x fewer
 * channels. */
int wsi_net_add_pool_bias(void* net, int in_id, int in_c_off, int C, int out_id, int out_c_off, int k, int stride,
                          int pad, int count_include_pad, const float* h_bias, int relu);
/* ReLU of the convolution added last only on its first `cols` output channels (multiple of 64): lets the bias-free,
 * ReLU-free 1x1 conv of a commuted pool branch ride in the same launch as its sibling 1x1 branch convs. */
int wsi_net_set_conv_relu_cols(void* net, int cols);
/* Residual of the convolution added last (added with residual_id = -1) = channels [res_c_off, res_c_off + Cout) of
 * tensor `residual_id`: the `out += identity` of a torchvision BasicBlock (TorchScript graph run by
 * run_inference.py:541) when `identity` is the downsample branch stored next to the block's first conv. */
int wsi_net_set_conv_residual_slice(void* net, int residual_id, int res_c_off);
/* Output channels >= cols of the convolution added last (odd kernel, 'same' padding) have weights at the centre tap
 * only: a 1x1 / stride-s downsample convolution stacked under the 3x3 / stride-s / pad-1 convolution that reads the
 * same block input.  Those channels get their own N tiles, for which only the centre tap is multiplied, so one launch
 * reads the block input once and writes both branches (three launches fewer per ResNet34 forward). */
int wsi_net_set_conv_center_cols(void* net, int cols);
int wsi_net_add_head(void* net, int in_id, const float* h_fc_w, const float* h_fc_b, int K);
int wsi_net_finalize(void* net);
/* d_input: the external input tensor (bf16 NHWC, layout of wsi_preproc_run); d_probs: f32 [B][K]
 * (softmax over K when K > 1, sigmoid when K == 1, run_inference.py:548-551). */
int wsi_net_forward(void* net, const void* d_input, int B, float* d_probs, void* stream);
int wsi_net_encode_input_map(void* net, int op_index, const void* d_base);
int wsi_net_tensor_ptr(void* net, int id, void** d_ptr, int* h_hwc3);
/* Copy `batch` images between a dense bf16 [batch][H][W][C] device buffer and tensor `id`; the
 * library may store a tensor with a one-pixel zero border (tensors touched by the flat 3x3 kernel),
 * so raw wsi_net_tensor_ptr access is only meaningful through these two. */
int wsi_net_read_tensor(void* net, int id, int batch, void* d_dense_out, void* stream);
int wsi_net_write_tensor(void* net, int id, int batch, const void* d_dense_in, void* stream);
/* 1 if op `op_index` is a conv that runs on the flat-slab kernel */
int wsi_net_op_is_flat(void* net, int op_index);
int wsi_net_num_launches(void* net);   /* kernels one forward launches */
int wsi_net_num_ops(void* net);        /* ops of the layer graph (>= launches: fused ops launch nothing) */
/* geometry (H, W, C) of the external input tensor the preprocess kernel must write; K of the head */
int wsi_net_input_geometry(void* net, int* h_hwc3);
int wsi_net_num_classes(void* net);
/* h_info4[0] = layout of the external input: 0 = bf16 [B][H][W][4] (geometry of wsi_net_input_geometry),
 * 1 = two row-parity planes of pixel pairs, bf16 [2][max_batch][P][P][8] (7x7/2 stems with 64 filters);
 * h_info4[1] = P; h_info4[2], h_info4[3] = low / high 32 bits of the plane stride in pixels.
 * wsi_net_input_bytes = bytes to allocate for that buffer (max_batch images + read slack). */
int wsi_net_input_layout(void* net, int* h_info4);
size_t wsi_net_input_bytes(void* net);
/* Per-op CUDA-event timing on the launching stream (bench roofline).  `slots` = number of most
 * recent forwards kept (0 disables).  wsi_net_profile_read synchronises the device and returns the
 * summed duration (ms) of each op over the recorded forwards, the op's flops per image and its kind
 * (0 conv, 1 max-pool, 2 avg-pool, 3 head). */
int wsi_net_set_profiling(void* net, int slots);
int wsi_net_profile_read(void* net, double* h_op_ms, double* h_op_flops, int* h_op_kind, int max_ops,
                         int* h_forwards);
double wsi_net_flops_per_image(void* net);
/* Algorithmic HBM bytes per image of the convolution launches (input slice + residual read once, output written
 * once; what roofline.traffic from ncu is compared with). */
double wsi_net_conv_bytes_per_image(void* net);
void wsi_net_destroy(void* net);

/* ---- host-buffer entry of the classification loop ------------------------------------------ */
/* Replaces one iteration of `for batch_imgs, batch_coords in loader:` in
 * wsinsight/modellib/run_inference.py:539-556 when the pixels are in HOST memory (what
 * read_region returns, wsinsight/modellib/data.py:276-280): h_rgba holds n RGBA u8 patches of
 * patch_size x patch_size pixels, patch i at h_rgba + i*patch_size*patch_pitch_bytes, rows
 * patch_pitch_bytes apart (multiple of 16).  submit() enqueues H2D copy -> fused preprocess ->
 * network forward -> D2H copy of the n x K probabilities into h_probs and returns at once; up to
 * `depth` batches are in flight (copy of batch i+1 overlaps the kernels of batch i).  h_rgba must
 * stay valid until the next sync(); h_probs is valid after sync().  Use pinned host memory
 * (wsi_host_alloc_pinned) or the copies degrade to staged synchronous ones. */
int wsi_hostpipe_create(void* net, int patch_size, long long patch_pitch_bytes, int out_size,
                        const float* mean, const float* stdv, int max_batch, int depth,
                        void** pipe_out);
/* Host buffers hold dense RGB patches u8 [n][ps][ps][3] (what read_region(...).convert("RGB") yields,
 * wsinsight/modellib/data.py:276-281) instead of RGBA rows: 25 % fewer bytes over PCIe, expanded to the RGBA stack on
 * the device.  channels 3 | 4 (default 4); call before the first submit. */
int wsi_hostpipe_set_host_channels(void* pipe, int channels);
int wsi_hostpipe_submit(void* pipe, const uint8_t* h_rgba, int n, float* h_probs);
int wsi_hostpipe_sync(void* pipe);
/* cudaStream_t the kernels of the pipe run on (for event timing by the caller) */
void* wsi_hostpipe_compute_stream(void* pipe);
/* bytes copied host->device / device->host since creation */
int wsi_hostpipe_bytes(void* pipe, long long* h2d_bytes, long long* d2h_bytes);
void wsi_hostpipe_destroy(void* pipe);
int wsi_host_alloc_pinned(size_t bytes, void** h_ptr);
int wsi_host_free_pinned(void* h_ptr);

/* ---- utilities -------------------------------------------------------------------------- */
/* Synthetic device-resident slide (bench/test input generator): fills an RGBA u8
 * image from a counter-based hash: smooth value-noise "tissue" field, H&E-like colours on a
 * near-white background, per-pixel noise.  The same function evaluated on the host is
 * wsi_synth_slide_host (used to regenerate sampled patches for the CPU baseline). */
int wsi_synth_slide(uint8_t* d_rgba, long long W, long long H, long long pitch_bytes,
                    unsigned seed, void* stream);
int wsi_synth_slide_host(uint8_t* h_rgba, long long x0, long long y0, int w, int h,
                         long long W, long long H, unsigned seed);
/* Box-average thumbnail of an RGBA slide -> RGB u8 Ht x Wt (bench input for K1; PIL-LANCZOS
 * thumbnail parity is a "next" row). */
int wsi_box_thumbnail(const uint8_t* d_rgba, long long W, long long H, long long pitch_bytes,
                      uint8_t* d_rgb, int Ht, int Wt, void* stream);

/* N1: the thumbnail the reference segments, bit-exact with Pillow.  Replaces
 * wsinsight/patchlib/pipeline.py:102-105 slide.get_thumbnail(thumbsize) -> tiffslide get_thumbnail ->
 * Image.thumbnail(size, LANCZOS) (reducing_gap = 2: Reduce.c box reduction + Resample.c Lanczos-3).
 * wsi_thumbnail_geometry: output size (aspect rule) and the integer reduce factors Pillow picks.
 * wsi_lanczos_thumbnail: RGBA u8 slide (alpha opaque) -> RGB u8 [out_h][out_w][3]; d_workspace of at least
 * wsi_thumbnail_workspace_bytes().  A reduce factor of 1 on exactly one axis is WSI_ERR_ARG (unsupported). */
int wsi_thumbnail_geometry(long long W, long long H, int max_w, int max_h, int* out_w, int* out_h, int* fx, int* fy);
size_t wsi_thumbnail_workspace_bytes(long long W, long long H, int max_w, int max_h);
int wsi_lanczos_thumbnail(const uint8_t* d_slide_rgba, long long W, long long H, long long pitch_bytes, int max_w,
                          int max_h, uint8_t* d_out_rgb, void* d_workspace, size_t workspace_bytes, void* stream);

/* K5 (BASELINE config 5): per-class density grid.  No upstream implementation exists;
 * the object centre follows wsinsight/insightlib/insight_helpers.py:26-27 (np.rint(minx + width / 2)), the class is
 * the first arg-max of the K probabilities.  ACCUMULATES into d_grid int32 [K][Gy][Gx]; objects outside the grid are
 * counted in *d_dropped.  Grids of several ranks are summed by the caller (one NCCL all-reduce). */
int wsi_density_accumulate(const int32_t* d_xywh, const float* d_probs, int n, int K, int bin_px, int Gx, int Gy,
                           int32_t* d_grid, int32_t* d_dropped, void* stream);

/* N4: the optional /images cache of the patch file.  wsi_extract_patches_rgb replaces
 * wsinsight/patchlib/io.py:19-48 (read_region + convert("RGB") per patch; zero outside the slide):
 * d_out_rgb u8 [n][ps][ps][3].  wsi_pack_patches_rgba turns such cached patches back into the RGBA stack
 * ([n * ps][pitch], A = 255) that wsi_preproc_plan_create accepts as a "slide" of width ps and height n * ps with
 * patch i at (0, i * ps) -- the device form of wsinsight/modellib/data.py:243-268. */
int wsi_extract_patches_rgb(const uint8_t* d_slide_rgba, long long W, long long H, long long pitch_bytes,
                            const int32_t* d_coords_xy, int n, int patch_size, uint8_t* d_out_rgb, void* stream);
int wsi_pack_patches_rgba(const uint8_t* d_rgb, int n, int patch_size, long long pitch_bytes, uint8_t* d_out_rgba,
                          void* stream);
/* Same packing at memory speed for patches [first_patch, first_patch + n) of the stack (shared-memory staged, aligned
 * vector loads): d_rgb 16-byte aligned with >= 16 readable bytes after its end; used by the host pipe's RGB mode. */
int wsi_pack_patches_rgba_fast(const uint8_t* d_rgb, int first_patch, int n, int patch_size, long long pitch_bytes,
                               uint8_t* d_out_rgba, void* stream);

/* N4: Macenko stain normalisation of RGB patches, per pixel (HistomicsTK deconvolution_based_normalization as the
 * reference calls it, wsinsight/modellib/data.py:293-300; matrix estimation wsinsight/modellib/run_inference.py:232-264).
 * h_M9 = W_target . inv(normalised, complemented W_source), row major, float64 (host).  d_rgb u8 [n_pixels][3];
 * in-place allowed.  HistomicsTK is not available offline: parity unpinned (oracle/stain.py restates it). */
int wsi_stain_normalize_rgb(const uint8_t* d_rgb_in, uint8_t* d_rgb_out, long long n_pixels, const double* h_M9,
                            void* stream);

/* a13 / N3: batch step of the cell-model tile stitcher, fused.  Replaces
 * wsinsight/modellib/tilefuse.py:219-278 TileRemapStitcher.accumulate_batch_torch (softmax, bilinear
 * align_corners=False resize h -> S, hv * alpha, tp renormalisation, D2H and the Python paste loop) with one kernel
 * writing into device canvases np_map f32 [H][W], hv_map f32 [H][W][2], tp_map f32 [H][W][K].
 * Inputs are the model outputs, f32 NCHW: np_logits [B][2][h][h], hv [B][2][h][h], tp_logits [B][K][h][h];
 * d_coords int32 [B][coord_stride] holds (minx, miny, ...) of each tile; its S x S interior lands at
 * (minx + halo, miny + halo), clipped to the canvas.  Tiles of one call must not overlap. */
int wsi_tilefuse_accumulate(const float* d_np_logits, const float* d_hv, const float* d_tp_logits,
                            const int32_t* d_coords, int coord_stride, int B, int K, int h, int S, int halo, float alpha,
                            int canvas_W, int canvas_H, float* d_np_map, float* d_hv_map, float* d_tp_map, void* stream);

/* N3 / a13: one padded tile of TileRemapStitcher.finalize (wsinsight/modellib/tilefuse.py:280-415): _proc_np_hv
 * (:39-103: np >= 0.5, label + remove_small_objects, min-max normalised HV maps, Sobel ksize 21, ridge / marker,
 * binary_fill_holes, 5x5 elliptical opening, label + remove_small_objects, GaussianBlur 3x3, marker-controlled
 * watershed) and _stitching_worker (:109-174: interior crop, bounding box, class-probability means, outer contour)
 * entirely on the device.
 *   d_np / d_hv / d_tp point at the first pixel of the PADDED tile inside the canvases (f32; hv 2 and tp K interleaved
 *   values per pixel), *_pitch = elements per canvas row.  [iy0, iy1) x [ix0, ix1) is the interior slice inside the
 *   padded H x W tile, (gx0, gy0) its slide coordinates.
 * Outputs (device): d_counts int32[4] = {instances, polygon points, error (1: > max_instances, 2: > max_poly_points),
 * markers}; d_bbox int32 [max_instances][4] (x, y, w, h in slide pixels, instances in label order == the order of
 * ndi.label); d_cls_mean f32 [max_instances][K]; d_poly_npts / d_poly_off int32 [max_instances] (0 points: the
 * reference drops polygons with < 3 vertices); d_poly_xy int32 [max_poly_points][2]; d_inst_map (optional) int32
 * [H][W] instance labels of the padded tile.  Differences to the reference are confined to exact floating-point ties
 * of the flooding order and to instances that the interior crop splits into several 8-connected pieces (the piece
 * holding the first raster pixel is traced; the reference takes the piece with the largest contour area). */
size_t wsi_cellpost_workspace_bytes(int H, int W, int K, int max_instances);
int wsi_cellpost_tile(const float* d_np, long long np_pitch, const float* d_hv, long long hv_pitch, const float* d_tp,
                      long long tp_pitch, int K, int H, int W, int iy0, int iy1, int ix0, int ix1, int gx0, int gy0,
                      int min_object_size, int max_instances, int max_poly_points, int* d_counts, int* d_bbox,
                      float* d_cls_mean, int* d_poly_npts, int* d_poly_off, int* d_poly_xy, int* d_inst_map, void* d_ws,
                      size_t ws_bytes, void* stream);

/* Connected-component labelling of one 0/1 image (the labelling stage of K1, exported for the cell path): min-index
 * union-find roots of both classes, 4-connected (conn8_fg: 8-connected foreground; frame: border background tied to
 * node H*W); area_cls in {0, 1}: pixel count of that class at every root (-1: none).  *d_L_out / *d_area_out point
 * into d_ws (wsi_seg_workspace_bytes(H, W) bytes). */
int wsi_ccl_label_u8(const uint8_t* d_mask, int H, int W, int conn8_fg, int frame, int area_cls, int** d_L_out,
                     int** d_area_out, void* d_ws, size_t ws_bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* WSI_B200_H_ */
