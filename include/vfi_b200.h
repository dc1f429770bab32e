/* vfi_b200.h -- C ABI of libvfi_b200.so: the B200 (sm_100a) replacement for the
 * frame-interpolation hot path of the gaming-VFI benchmark study.
 *
 * Every entry point names the reference interface it replaces (paths relative to the
 * reference tree).  All functions return 0 on success or a negative VFI_E_* code;
 * vfi_last_error(h) gives the message.  Unless stated otherwise all buffers are DEVICE
 * pointers owned by the caller; the library owns only the handle's weights + workspace.
 * `stream` is a cudaStream_t passed as void* (0 = legacy default stream).  A handle is
 * not thread-safe; different handles are independent (the reference is single-threaded:
 * scripts/run_benchmarks.py:142 hot loop).  All calls on one handle share its workspace, so the
 * library ORDERS them: a call issued on a different stream than the previous call first waits
 * (on the device, cudaStreamWaitEvent) for that previous call to finish.  Calls from several
 * streams are therefore safe but run one after the other; use one handle per stream to overlap.
 * Every entry point leaves the calling thread's current CUDA device as it found it.
 * There is no CPU fallback: every call fails with VFI_E_CUDA when no sm_100 device exists.
 */
#ifndef VFI_B200_H
#define VFI_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct vfi_handle vfi_handle;

#define VFI_OK 0
#define VFI_E_ARG (-1)     /* bad argument (shape, null pointer, unsupported scale ...) */
#define VFI_E_CUDA (-2)    /* CUDA runtime / driver error, message in vfi_last_error   */
#define VFI_E_STATE (-3)   /* weights not loaded                                         */
#define VFI_E_NOMEM (-4)

#define VFI_PRECISION_FP32 0 /* CUDA-core fp32 convs: 1e-4 parity mode                   */
#define VFI_PRECISION_BF16 1 /* tcgen05 bf16 implicit-GEMM convs, fp32 accumulate/flow   */

#define VFI_ROUTE_RIFE 0
#define VFI_ROUTE_VFIMAMBA 1
#define VFI_ROUTE_SCENE_CHANGE 2

#define VFI_IFNET_PARAMS 5664776

/* lifetime ------------------------------------------------------------------------- */
int vfi_create(int device, vfi_handle** out);
void vfi_destroy(vfi_handle* h);
const char* vfi_last_error(vfi_handle* h);
/* Stream-ordering fence: makes `stream` wait for the handle's previous call (if that ran on another stream) and records it as
 * the handle's latest user; launches nothing.  Calls captured into a CUDA graph (cudaStreamBeginCapture on `stream` around
 * vfi_ifnet_infer: the single-pair latency path of RIFEModel.interpolate) skip this bookkeeping, so the graph's owner
 * brackets every replay with vfi_fence before and after. */
int vfi_fence(vfi_handle* h, void* stream);

/* number of kernels this handle has launched since creation (bench.py "gpu_launches") */
long long vfi_launch_count(vfi_handle* h);

/* device memory the handle owns right now: workspace arena + reduction scratch + packed weights (bytes).  PyTorch's
 * allocator statistics do not see these; the host wrappers add it to InferenceResult.vram_peak_mb (models/base.py:140-158) */
size_t vfi_device_bytes(vfi_handle* h);

/* Tuning / test knobs (defaults are the fastest measured setting; every setting computes the same result):
 *   "conv_v2"       1: stride-1 3x3 convs whose weights fit in shared memory run on the resident-weight halo-tile
 *                      kernel (csrc/conv_v2.cu) and the stride-2 ones on csrc/conv_s2.cu; 0: TMA-per-tap kernel.
 *   "v2_tma"        1: conv_v2 halo tiles arrive by TMA tensor boxes; 0: by cp.async producer warps.
 *   "fuse_gather"   1: the finest block's input gather is fused with conv0[0] (and hands the summed flow to the blend).
 *   "v2_pm"         1: c = 64 layers use a pixel-major SWIZZLE_128B halo tile (one 128-byte-row TMA box); 0: chunk-major tile (16-byte rows).
 *   "v2_res_global" 0: ResConv residual from the shared-memory tile; 1: from global memory.
 *   "l2_chunk_mb"   0: off; n: run each ResConv chain depth-first over item groups of n MB of activations.
 *   "gc_persistent" 1: the fused finest-block gather is persistent (staging of tile i+1 overlaps tile i's MMAs / epilogue); 0: one tile per CTA.
 *   "fuse_gather3"  0: block 3 runs gather_s2 + conv_s2; 1: its gather is fused with its conv0[0] like the finest block's
 *                      (same results to the bf16 tolerance; measured 3 % slower per step on B200, kept as a variant).
 *   "t_fast"        1: the gathers and the blend walk their tiles timestep-fastest (the T items of a pair side by side, so the pair's
 *                      images are read from DRAM once, not once per timestep); 0: item-major order; 2: as 1 with the pairs next
 *                      in the fused finest gather (timestep, pair, x, y: measured equal to 1).  Bit-identical results.
 *   "gc_ablate": debug only (phase ablation of the fused finest gather).
 *   "v2_ablate", "v2_trace": debug only (tools/bench_conv.py, tools/v2_trace.py).
 *   "ws_guard"      debug only: 1 makes vfi_ifnet_infer_ex and vfi_farneback fill their workspace with a byte pattern, leave a 4 KiB gap after every
 *                      intermediate buffer they carve from it, and fail with VFI_E_CUDA if a kernel wrote into a gap (one extra
 *                      memset, one check kernel and a stream synchronise per call, so not under stream capture; 2 = the check's self-test: it
 *                      overwrites one gap byte itself and the call must fail; tests/test_gpu_bounds.py). */
int vfi_set_option(vfi_handle* h, const char* key, int value);

/* IFNet weights.  Replaces Model.load_model(dir, -1) (models/sota/rife_wrapper.py:102-103,
 * scripts/run_experiments.py:266).  `params` is a HOST array of VFI_IFNET_PARAMS floats:
 * the v4.25 IFNet state_dict tensors flattened (PyTorch layouts: Conv2d [Co,Ci,kh,kw],
 * ConvTranspose2d [Ci,Co,kh,kw]) in this order:
 *   encode.cnn0.{weight,bias} cnn1 cnn2 cnn3,
 *   for b in block0..block4: conv0.0.0.{weight,bias}, conv0.1.0.{weight,bias},
 *       for j in 0..7: convblock.j.conv.{weight,bias}, convblock.j.beta,
 *       lastconv.0.{weight,bias}
 * `lrelu_slope` NULL => LeakyReLU(0.2) (v4.25); otherwise one slope for every activation. */
int vfi_ifnet_load_weights(vfi_handle* h, const float* params, size_t n_params, const float* lrelu_slope);

/* Replaces RIFEModel.interpolate's device work (rife_wrapper.py:127-152): to_tensor
 * (models/base.py:175-186), pad_to_multiple reflect (base.py:203-225), Model.inference(
 * img0,img1,timestep=t,scale) for each t (rife_wrapper.py:141-146), unpad (base.py:227)
 * and to_numpy's *255/clip/truncate (base.py:188-201).
 *   img0,img1 : B x H x W x 3 uint8 RGB          timesteps : HOST array of T floats
 *   scale     : 1.0, 0.5 or 0.25 (rife_wrapper.py:144; run_experiments.py:192,214)
 *   hud_mask  : nullable, B x H x W uint8; where non-zero the output pixel is copied from
 *               img0 (t < 0.5) or img1 (adaptive_pipeline.py:399-409)
 *   out       : B x T x H x W x 3 uint8          out_f32 : nullable, same shape, float [0,1] unclamped
 *   flow_out  : nullable, B x T x H x W x 4 float final flow (debug / parity)             */
int vfi_ifnet_infer(vfi_handle* h, const uint8_t* img0, const uint8_t* img1, int B, int H, int W,
                    const float* timesteps, int T, float scale, int precision,
                    const uint8_t* hud_mask, uint8_t* out, float* out_f32, float* flow_out, void* stream);

/* The same with the conventions of the reference's experiment runner, scripts/run_experiments.py:365-386
 * (rife_interpolate; duplicated at gui/benchmark_proper.py:90-101):
 *   VFI_FLAG_BGR              frames in and out are BGR (cv2 order); the network sees RGB (:381-382 flip, :386 flip back)
 *   VFI_FLAG_PAD_REFLECT_EDGE the bottom/right padding is cv2.BORDER_REFLECT (border pixel repeated, :379-380) instead
 *                             of the reflect-101 of models/base.py:203-225, and the
 *                             padding multiple is 128 at every scale, as :373-375 computes it
 * flags = 0 is vfi_ifnet_infer. */
#define VFI_FLAG_BGR 1
#define VFI_FLAG_PAD_REFLECT_EDGE 2
int vfi_ifnet_infer_ex(vfi_handle* h, const uint8_t* img0, const uint8_t* img1, int B, int H, int W,
                       const float* timesteps, int T, float scale, int precision, int flags,
                       const uint8_t* hud_mask, uint8_t* out, float* out_f32, float* flow_out, void* stream);

/* The fused backward-warp + sigmoid-blend kernel on its own (IFNet.forward tail:
 * warp(img0,flow[:, :2]), warp(img1,flow[:, 2:4]), sigmoid(mask), blend -- reached through
 * rife_wrapper.py:146).  img0/img1: N x H x W x 3 uint8; flow: N x H x W x 4 float;
 * mask: N x H x W float (pre-sigmoid); out: N x H x W x 3 uint8; out_f32 nullable.      */
int vfi_warp_blend(vfi_handle* h, const uint8_t* img0, const uint8_t* img1, const float* flow,
                   const float* mask, int N, int H, int W, uint8_t* out, float* out_f32, void* stream);

/* compute_psnr_simple (evaluation/metrics.py:299-315) for n same-shape uint8 frame pairs of
 * `elems` bytes each.  sse_out: n x uint64 exact integer sum of squared differences;
 * psnr_out: n x double (inf when sse == 0).  Either may be NULL.                          */
int vfi_psnr_u8(vfi_handle* h, const uint8_t* a, const uint8_t* b, int n, size_t elems,
                unsigned long long* sse_out, double* psnr_out, void* stream);

/* The five integer moments behind the experiment runner's GLOBAL (single-window) metrics, scripts/run_experiments.py:292-312
 * calc_psnr / calc_ssim (every published number of the study; duplicated at gui/benchmark_proper.py:31-44): for n pairs of
 * `elems` uint8 values, out[5*i + 0..4] = sum a, sum b, sum a^2, sum b^2, sum a*b (int64, exact).  The float64 formulas
 * are finished on the host with the reference's own expressions.                                                        */
int vfi_moments_u8(vfi_handle* h, const uint8_t* a, const uint8_t* b, int n, size_t elems, long long* out, void* stream);

/* NO COUNTERPART IN THE REFERENCE: added for callers that may not call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same inputs, outputs and results as
 * vfi_psnr_u8, but the per-pair SSE accumulators live in a static __device__ array inside the library instead of the
 * handle's cudaMalloc'd scratch: the call allocates nothing (no cudaMalloc / cudaFree / allocating library routine).
 *   limit  : n <= VFI_NOALLOC_MAX_PAIRS frame pairs per call (32 KB of device memory held by every GPU that loads the
 *            library, until exit); larger n is rejected with VFI_E_ARG -- split the call or use vfi_psnr_u8.
 *   sharing: every call on a device uses the same array, which keeps the previous call's sums; each call clears the entries
 *            it uses on `stream` first.  Calls that overlap on the GPU (different streams without ordering) race: serialise them.
 *   set-up : the array's device address is looked up once per device on the first call (make that call outside a capture). */
#define VFI_NOALLOC_MAX_PAIRS 4096
int vfi_psnr_u8_noalloc(vfi_handle* h, const uint8_t* a, const uint8_t* b, int n, size_t elems,
                        unsigned long long* sse_out, double* psnr_out, void* stream);

/* compute_ssim_simple (evaluation/metrics.py:318-335; skimage structural_similarity on the
 * cv2 gray images, 7x7 uniform window) when downscale == 1; AdaptiveRouter.detect_scene_change's
 * score (adaptive_pipeline.py:140-160: gray, x0.25 INTER_LINEAR, SSIM) when downscale == 4.
 * a,b: n x H x W x 3 uint8 RGB; out: n x double.                                           */
int vfi_ssim_gray_u8(vfi_handle* h, const uint8_t* a, const uint8_t* b, int n, int H, int W,
                     int downscale, double* out, void* stream);

/* compute_psnr_simple + compute_ssim_simple of the same n frame pairs from ONE read of the frames (the evaluation loop of
 * scripts/run_benchmarks.py:174-178 calls both per frame; evaluation/metrics.py:299-335).  a,b: n x H x W x 3 uint8 RGB.
 * sse_out (n x uint64, exact) and psnr_out (n x double) may be NULL; ssim_out: n x double.                              */
int vfi_psnr_ssim_u8(vfi_handle* h, const uint8_t* a, const uint8_t* b, int n, int H, int W,
                     unsigned long long* sse_out, double* psnr_out, double* ssim_out, void* stream);

/* AdaptiveRouter.compute_motion's reductions (adaptive_pipeline.py:130-138) over a dense flow
 * field n x H x W x 2 float: out[3*i+0..2] = mean, max, std (population) of |flow|.        */
int vfi_flow_stats(vfi_handle* h, const float* flow, int n, int H, int W, double* out, void* stream);

/* Dense Farneback optical flow on the device -- replaces cv2.calcOpticalFlowFarneback(prev, next, None, pyr_scale,
 * levels, winsize, iterations, poly_n, poly_sigma, flags=0) at models/novel/adaptive_pipeline.py:119-128
 * (0.5, 3, 15, 3, 5, 1.2; same call at evaluation/metrics.py:379-381) and scripts/run_experiments.py:359-362,460
 * (0.5, 1, 15, 1, 5, 1.1).  OpenCV's algorithm (optflowgf.cpp), restated in oracle/farneback_ref.py.
 *   img0,img1 : n x H x W x channels uint8; channels 3 = RGB (converted with cv2's RGB2GRAY integers), 1 = gray
 *   flow_out  : nullable, n x H x W x 2 float (u, v)
 *   stats_out : nullable, n x 3 double = mean, max, std of |flow| (compute_motion, adaptive_pipeline.py:130-138);
 *               the flow never leaves the device.  At least one of the two outputs must be given.
 * Limits: 0 < pyr_scale < 1, poly_n 5 or 7, winsize <= 31, flags 0 (box window, no initial flow).                */
int vfi_farneback(vfi_handle* h, const uint8_t* img0, const uint8_t* img1, int n, int H, int W, int channels,
                  double pyr_scale, int levels, int winsize, int iterations, int poly_n, double poly_sigma,
                  float* flow_out, double* stats_out, void* stream);

/* The experiment runner's motion-threshold router, scripts/run_experiments.py:848-874 (the "adaptive" method every
 * adaptive_* result of the study used): cv2.resize(gray, (480, 270)) of both frames -> vfi_farneback(0.5, 1, 15, 1, 5, 1.1)
 * -> mean |flow| > motion_thresh ? RIFE : cv2.addWeighted(prev, 1-t, curr, t, 0).
 *   vfi_resize_gray_u8 : cv2.resize(cv2.cvtColor(src, COLOR_{RGB,BGR}2GRAY), (dw, dh)), INTER_LINEAR uint8 fixed point,
 *                        bit-exact.  src: n x H x W x channels (3: RGB, or BGR when bgr != 0; 1: already gray); dst: n x dh x dw.
 *   vfi_add_weighted_u8: cv2.addWeighted(a, alpha, b, beta, 0) on uint8 (float multiply-add, round half to even,
 *                        saturate); bit-exact whenever a*alpha and b*beta are exact in float (t = k/4, as the runner uses). */
int vfi_resize_gray_u8(vfi_handle* h, const uint8_t* src, int n, int H, int W, int channels, int bgr, int dh, int dw,
                       uint8_t* dst, void* stream);
int vfi_add_weighted_u8(vfi_handle* h, const uint8_t* a, const uint8_t* b, size_t elems, double alpha, double beta,
                        uint8_t* out, void* stream);

/* cv2.Canny(gray, low, high) (aperture 3, L1 gradient) -- the edge-density term of the MCAR complexity score,
 * scripts/run_experiments.py:466-468 (`np.mean(edges > 0)`).  gray: n x H x W uint8.  edges_out: nullable, n x H x W uint8
 * (255 / 0, bit-exact with OpenCV); density_out: nullable, n x double = edge pixels / (H*W).  Synchronises `stream`
 * (the hysteresis runs to its fixed point under host control).                                                       */
int vfi_canny_u8(vfi_handle* h, const uint8_t* gray, int n, int H, int W, double low, double high,
                 uint8_t* edges_out, double* density_out, void* stream);

/* Optional planar YUV 4:2:0 (I420) RESULT format: cv2.cvtColor(frame, COLOR_RGB2YUV_I420) (COLOR_BGR2YUV_I420 when bgr != 0),
 * bit-exact (BT.601 limited range, chroma of the top-left pixel of each 2 x 2 block).  The experiment runner hands every
 * interpolated frame to a libx264 encoder (scripts/run_experiments.py:861-868 `ff_proc.stdin.write(mid.tobytes())`), which
 * converts to yuv420p; converting on the device halves the device -> host bytes of an interpolated frame (1.5 B/px, not 3).
 *   src: n x H x W x 3 uint8 (H, W even); dst: n x (H*3/2) x W uint8 = per frame Y (H x W), U (H/2 x W/2), V (H/2 x W/2). */
int vfi_rgb_to_yuv420_u8(vfi_handle* h, const uint8_t* src, int n, int H, int W, int bgr, uint8_t* dst, void* stream);

/* MCAR tier 2, scripts/run_experiments.py:498-503 / :526-531: cv2.remap(src, x + flow_x * t, y + flow_y * t, cv2.INTER_LINEAR)
 * for a uint8 image (H x W x channels, BORDER_CONSTANT 0) and a dense flow field (H x W x 2 float, e.g. vfi_farneback's
 * flow_out); OpenCV's 5-bit fixed-point bilinear path, bit-exact.  Follow with vfi_add_weighted_u8(dst, 1-t, curr, t). */
int vfi_remap_flow_u8(vfi_handle* h, const uint8_t* src, const float* flow, double t, int H, int W, int channels,
                      uint8_t* dst, void* stream);

/* detect_particles' frequency term (adaptive_pipeline.py:180-183): variance of the 3x3
 * Laplacian (reflect-101) of the gray image.  rgb: n x H x W x 3; out: n x double.        */
int vfi_laplacian_var(vfi_handle* h, const uint8_t* rgb, int n, int H, int W, double* out, void* stream);

/* AdaptiveRouter.detect_hud (adaptive_pipeline.py:193-235) for one frame.  hist_ring:
 * hist_len x 180 x 320 uint8 ring of gray thumbnails; *count (HOST, in/out) = frames pushed so
 * far.  mask_out: H x W uint8 (0/1); coverage_out: 1 double (device).  With fewer than 5 frames
 * the mask is all zero and coverage 0, as in the reference.                                */
int vfi_hud_update(vfi_handle* h, const uint8_t* rgb, int H, int W, uint8_t* hist_ring, int hist_len,
                   long long* count, double var_threshold, uint8_t* mask_out, double* coverage_out,
                   void* stream);

/* The decision tree of AdaptiveRouter.analyze (adaptive_pipeline.py:242-275) on HOST doubles.
 * Returns the route in *route (VFI_ROUTE_*), confidence in *confidence, particle flag/score. */
int vfi_route_decide(double ssim_score, double motion_mean, double motion_max, double motion_std,
                     double laplacian_var, double scene_thr, double particle_thr, double motion_lo,
                     double motion_hi, int* route, double* confidence, int* has_particles,
                     double* particle_score);

/* Unit-test hook: run ONE conv layer of the loaded IFNet on caller data in both precisions'
 * layouts.  layer ids: csrc/common.cuh (VFI_NUM_LAYERS table) / ifnet_layers.py.  x: N x Hin x Win x Cin_pad (fp32 NHWC),
 * y: N x Hout x Wout x Cout_pad fp32 NHWC out (for lastconv: N x 4Hin x 4Win x 16).
 * residual (ResConv layers only): NULL => the layer input itself, as in the net.          */
int vfi_debug_conv(vfi_handle* h, int layer, int precision, const float* x, int N, int Hin, int Win,
                   const float* residual, float* y, void* stream);

/* Per-kernel-class device timing for bench.py's roofline block.  While enabled every kernel launch is
 * bracketed by CUDA events on its stream; vfi_profile_read synchronises, returns summed milliseconds and
 * launch counts per class (0 pad, 1 head_cnn0, 2 conv_tcgen05, 3 conv_direct, 4 deconv_direct, 5 gather,
 * 6 gather_conv0a (finest block: gather fused with conv0[0]), 7 warp_blend, 8 reduce, 9 misc) and clears the log. */
#define VFI_PROFILE_CLASSES 10
int vfi_profile_enable(vfi_handle* h, int on);
int vfi_profile_read(vfi_handle* h, double* ms, long long* counts, int ncat);

#ifdef __cplusplus
}
#endif
#endif
