// common.cuh -- handle, layer table and helpers shared by every translation unit of libvfi_b200.so
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <string>
#include <vector>

#include "../../include/vfi_b200.h"

typedef __nv_bfloat16 bf16;

// ---- IFNet v4.25 layer table --------------------------------------
// ids: 0..3 = Head cnn0..cnn3; block b: base = 4 + 11*b: +0 conv0a, +1 conv0b, +2..+9 ResConv, +10 lastconv
#define VFI_NUM_LAYERS 59
enum LayerKind { LK_CONV3 = 0, LK_DECONV4 = 1 };

#define VFI_MAX_STEPS 32
// One k-step of the implicit GEMM: one (tap, channel-chunk) A tile and the B rows that consume it.
struct TcSeg {
  int16_t b_row_off;  // first row inside this step's B tile
  int16_t n;          // MMA N
  int16_t tmem_col;   // accumulator column offset
  int16_t accumulate; // 0 => first MMA into these columns (overwrite)
};
struct TcStep {
  int16_t dy, dx;    // A tile coordinate offset (in the A tensor's own pixel grid)
  int16_t plane;     // phase plane of a stride-2 input (0 for plain tensors)
  int16_t c0;        // first input channel of the chunk
  int32_t b_row;     // first row of this step's B tile in the packed weight matrix
  int16_t b_rows;    // rows to fetch
  int16_t nseg;
  TcSeg seg[2];
};

struct Layer {
  int kind, cin, cout, stride;
  int cin_pad;    // channel count of the NHWC input tensor (multiple of 16; 4 for cnn0)
  int cout_pad;   // channel stride of the output / padded weight columns (multiple of 4)
  bool has_beta, act;
  // fp32 path: conv [9][cin_pad][cout_pad]; deconv [16][cin_pad][cout_pad]
  float* w32 = nullptr;
  float* bias = nullptr;  // cout_pad floats (zero padded)
  float* beta = nullptr;  // cout floats or null
  // tcgen05 path
  bf16* wtc = nullptr;    // packed [rows_total][cblk] K-major
  int cblk = 0, rows_total = 0, nsteps = 0, n_total = 0, b_box = 0;
  TcStep steps[VFI_MAX_STEPS];
  float* bias_tc = nullptr;  // bias laid out by accumulator column (n_total floats)
  float h_bias_v2[128] = {0}, h_beta[128] = {0};  // host copies for conv_v2's constant-bank epilogue parameters
  float* bias_v2 = nullptr;  // conv_v2.cu Head.cnn3 instance: bias per (phase, channel) column
  bf16* ws2 = nullptr;       // conv_s2.cu resident-weight image of a stride-2 conv (null: not s2-capable)
  bf16* wv2 = nullptr;       // conv_v2.cu resident-weight image [tap][cin/8][cout][8] (null: layer not v2-capable)
};

struct vfi_handle {
  int device = 0;
  int sm_count = 148;
  std::string err;
  long long launches = 0;
  bool weights_loaded = false;
  float slope = 0.2f;
  Layer layers[VFI_NUM_LAYERS];
  float head0_w[27 * 16] = {0}, head0_b[16] = {0};  // Head.cnn0 weights [tap][ci][co] / 255 and bias: kernel-parameter operands
  // grow-only workspace arena
  void* ws = nullptr;
  size_t ws_bytes = 0;
  // small scratch for reductions
  void* red = nullptr;
  size_t red_bytes = 0;
  // One stream at a time may use the handle's scratch (ws, red, tdev).  Every entry point waits, on ITS stream, for the
  // event the previous entry point recorded when that one ran on a different stream, so calls from several streams
  // (ClipStreamer's compute stream next to a router call on the current stream) serialise instead of racing.
  cudaEvent_t scratch_ev = nullptr;
  cudaStream_t scratch_stream = nullptr;
  bool scratch_used = false;
  size_t weight_bytes = 0;      // device bytes held by the packed weights (vfi_device_bytes)
  unsigned long long* hud_cnt = nullptr;  // {mask pixel count, CTA tickets} of the fused HUD kernel; zero between launches
  float* tdev = nullptr;        // the call's timesteps on the device (64 floats); re-uploaded only when they change
  float t_host[64] = {0};
  int t_n = 0;
  void* tmap_encode = nullptr;  // cuTensorMapEncodeTiled entry point
  int opt_l2_chunk_mb = 0;      // vfi_set_option("l2_chunk_mb"): the ResConv chain of a block runs over groups of items whose
                                // activations (in + out) fit this many MB, so a layer reads what the previous one left in L2; 0 = off
                                // (measured slower at every size: +15 us of fill / drain per extra launch outweighs the L2 hits)
  int opt_v2_res_global = 0;    // vfi_set_option("v2_res_global"): ResConv residual read from global memory, not the smem tile (slower)
  int opt_v2_pm = 1;            // vfi_set_option("v2_pm"): pixel-major 128-byte-swizzled halo tiles for c = 64 (one wide TMA box per
                                // tile, tap-shifted SWIZZLE_128B descriptors).  Same speed as the chunk-major tile while the epilogue chain
                                // bound the layer (round 1); 1.5 % of the conv class faster since the two-group epilogue (round 2): default
  int opt_v2_tma = 1;           // vfi_set_option("v2_tma"): conv_v2 halo tiles by TMA (one thread) instead of cp.async producers
  int opt_fuse_gather = 1;      // vfi_set_option("fuse_gather"): finest block's gather fused with conv0[0]
  int opt_v2_trace = 0;         // debug only: conv_v2 dumps CTA 0's pipeline time stamps
  int opt_v2_ablate = 0;        // debug only: see V2Params::ablate
  int opt_ws_guard = 0;         // debug only: vfi_ifnet_infer_ex checks pattern gaps between its workspace buffers (abi.cu ws_guard_check)
  int opt_fuse_gather3 = 0;     // vfi_set_option("fuse_gather3"): block 3's (s = 2) gather fused with its conv0[0].  OFF: measured 3 % SLOWER
                                // per step than gather_s2 + conv_s2 (18 instead of 24 resident warps on a latency-bound gather)
  int opt_gc_persistent = 1;    // vfi_set_option("gc_persistent"): persistent, phase-overlapped form of the fused finest-block gather
  int opt_t_fast = 1;           // vfi_set_option("t_fast"): the gathers and the blend walk their tiles timestep-fastest, so the T items of a
                                // pair share the pair's image reads in L2 (item-major order re-read every image from DRAM once per timestep)
  int opt_gc_ablate = 0;        // debug only: gather_conv0a phases off (1 staging, 2 gather, 4 MMA + epilogue, 8 flow evaluation)
  int opt_conv_v2 = 1;          // vfi_set_option("conv_v2"): resident-weight halo-tile conv kernel where it applies
  // optional per-kernel-class timing (vfi_profile_enable): event pairs around every launch
  bool profile = false;
  std::vector<cudaEvent_t> prof_ev;   // begin,end,begin,end...
  std::vector<int> prof_cat;
};

enum ProfCat { PC_PAD = 0, PC_HEAD0, PC_CONV_TC, PC_CONV_DIRECT, PC_DECONV_DIRECT, PC_GATHER, PC_GATHER_CONV, PC_BLEND,
               PC_REDUCE, PC_MISC, PC_COUNT };
void prof_begin(vfi_handle* h, int cat, cudaStream_t s);
void prof_end(vfi_handle* h, cudaStream_t s);
struct ProfScope {
  vfi_handle* h; cudaStream_t s;
  ProfScope(vfi_handle* h_, int cat, cudaStream_t s_) : h(h_), s(s_) { if (h->profile) prof_begin(h, cat, s); }
  ~ProfScope() { if (h->profile) prof_end(h, s); }
};

// Entry-point scope: makes h->device current for the call and restores the caller's device afterwards (the library must
// not move the thread's CUDA device behind PyTorch's back), and orders this call after the handle's previous call when
// that one ran on another stream.
struct CallScope {
  vfi_handle* h; cudaStream_t s; int prev_dev = -1; bool ok = true, capturing = false;
  CallScope(vfi_handle* h_, cudaStream_t s_) : h(h_), s(s_) {
    if (cudaGetDevice(&prev_dev) != cudaSuccess) prev_dev = -1;
    if (prev_dev != h->device && cudaSetDevice(h->device) != cudaSuccess) { ok = false; return; }
    // a call being captured into a CUDA graph (RIFEModel's single-pair path) records no cross-stream bookkeeping: the
    // graph owner replays it on one stream and orders it against other users of the handle itself
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    capturing = cudaStreamIsCapturing(s, &cs) == cudaSuccess && cs != cudaStreamCaptureStatusNone;
    if (!capturing && h->scratch_used && h->scratch_stream != s) cudaStreamWaitEvent(s, h->scratch_ev, 0);
  }
  ~CallScope() {
    if (ok && !capturing) { cudaEventRecord(h->scratch_ev, s); h->scratch_stream = s; h->scratch_used = true; }
    if (prev_dev >= 0 && prev_dev != h->device) cudaSetDevice(prev_dev);
  }
};
#define VFI_ENTER(h, stream)                                                  \
  CallScope _scope((h), (cudaStream_t)(stream));                              \
  if (!_scope.ok) VFI_FAIL(h, VFI_E_CUDA, "cudaSetDevice(%d) failed", (h)->device)

#define VFI_FAIL(h, code, ...)                               \
  do {                                                       \
    char _b[512];                                            \
    snprintf(_b, sizeof(_b), __VA_ARGS__);                   \
    (h)->err = _b;                                           \
    return (code);                                           \
  } while (0)

#define VFI_CUDA(h, expr)                                                                   \
  do {                                                                                      \
    cudaError_t _e = (expr);                                                                \
    if (_e != cudaSuccess)                                                                  \
      VFI_FAIL(h, VFI_E_CUDA, "%s:%d %s: %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
  } while (0)

#define VFI_LAUNCHED(h)                                                                     \
  do {                                                                                      \
    (h)->launches++;                                                                        \
    cudaError_t _e = cudaGetLastError();                                                    \
    if (_e != cudaSuccess)                                                                  \
      VFI_FAIL(h, VFI_E_CUDA, "%s:%d launch: %s", __FILE__, __LINE__, cudaGetErrorString(_e)); \
  } while (0)

// One 32-byte store = one full DRAM/L2 sector per lane.  Two 16-byte stores of the same 32 bytes are two half-sector
// writes per lane (lanes sit >= 64 B apart in every NHWC epilogue here) and measured 2x slower on B200.
#ifdef __CUDACC__
__device__ __forceinline__ void st_v8(void* p, uint4 a, uint4 b) {
  asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(p), "r"(a.x), "r"(a.y), "r"(a.z),
               "r"(a.w), "r"(b.x), "r"(b.y), "r"(b.z), "r"(b.w)
               : "memory");
}
__device__ __forceinline__ void st_v8f(float* p, float4 a, float4 b) {
  asm volatile("st.global.v8.f32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(p), "f"(a.x), "f"(a.y), "f"(a.z),
               "f"(a.w), "f"(b.x), "f"(b.y), "f"(b.z), "f"(b.w)
               : "memory");
}
#endif

static inline int cdiv(long long a, long long b) { return (int)((a + b - 1) / b); }

// arena helpers (abi.cu)
int vfi_ws_reserve(vfi_handle* h, size_t bytes);
// Carves the buffers of one call out of the handle's workspace (first pass with base = nullptr sizes it).  With
// vfi_set_option("ws_guard") every buffer is followed by a pattern gap that vfi_ws_guard_check verifies after the call's kernels.
#define WS_GUARD_BYTE 0xA5
#define WS_GUARD_BYTES 4096
#define WS_GUARD_MAX_GAPS 64
struct Carve {
  uint8_t* base; size_t off;
  size_t guard = 0;                                      // vfi_set_option("ws_guard"): bytes of pattern kept after every buffer
  std::vector<unsigned long long>* gaps = nullptr;       // (offset, length) of each guard gap, filled on the carving pass
  template <typename T> T* take(size_t n) {
    off = (off + 1023) & ~(size_t)1023;
    T* p = base ? reinterpret_cast<T*>(base + off) : nullptr;
    off += n * sizeof(T);
    if (guard) {
      // the gap runs from the buffer's last byte to the next buffer's 1 KiB-aligned start
      const size_t end = (off + guard + 1023) & ~(size_t)1023;
      if (gaps && base) { gaps->push_back(off); gaps->push_back(end - off); }
      off = end;
    }
    return p;
  }
};
// device scratch of the check: 2 counters + the gap list; take it with cv.guard = 0, outside the guarded range
inline unsigned long long* ws_guard_scratch(Carve& cv) {
  const size_t g = cv.guard;
  cv.guard = 0;
  unsigned long long* p = cv.take<unsigned long long>(2 + 2 * WS_GUARD_MAX_GAPS);
  cv.guard = g;
  return p;
}
int vfi_ws_guard_check(vfi_handle* h, const std::vector<unsigned long long>& gaps, unsigned long long* dev, cudaStream_t s);
int vfi_red_reserve(vfi_handle* h, size_t bytes);

// ---- elementwise.cu -----------------------------------------------------------------------
int k_pad_rgbx(vfi_handle* h, const uint8_t* src, uint8_t* dst, int B, int H, int W, int Hp, int Wp, cudaStream_t s, int edge = 0, int bgr = 0);
int k_head_cnn0(vfi_handle* h, const Layer& L, const uint8_t* imgs, void* out, int NI, int Hp, int Wp, bool bf, cudaStream_t s);
// The lastconv outputs of the blocks run so far.  low[k] = [N, Hp/s[k], Wp/s[k], 4*c4[k]] fp32 (flow 4, mask 1,
// feat 8).  The reference's full-resolution flow / mask / feat tensors (IFBlock.forward's F.interpolate(scale) and
// IFNet.forward's flow = flow + fd) are never materialised: a consumer evaluates, at the pixels it needs,
//   flow = sum_k s[k] * bilerp(low[k][0:4]),  (mask, feat) = bilerp(low[n-1][4:13])        
// one axis of F.interpolate(bilinear, align_corners=False) from a grid s times coarser: v = h * src[i0] + l * src[i1]
struct __align__(16) Tap {
  int i0, i1;
  float h, l;
};
// bf16-mode storage of a coarse level (blocks 0..3), one array per field so a gather touches only what it needs:
// [N,Hl,Wl] x { flow delta: 4 fp32 | mask + feat 0..6: 8 bf16 | feat 7: 1 bf16 }.  The flow stays fp32 (it is summed over
// blocks and addresses pixels); mask / feat feed the next block's bf16 convolution input anyway.
struct LowPlanes {
  float* flow;
  bf16* mf;
  bf16* f7;
};
struct LowSet {
  const float* p[5];
  const bf16* mf[5];  // non-null: the level is stored as LowPlanes (p = flow plane, c4 = 1)
  const bf16* f7[5];
  const Tap* xt[5];  // per level (optional): taps of every padded-frame column [Wp] and row [Hp], see k_build_taps
  const Tap* yt[5];
  int s[5];
  int c4[5];  // float4 per texel: 4 (16 channels) or 2 (the final block's compact flow+mask)
  int Hl[5], Wl[5];  // Hp / s, Wp / s
  float rs[5], fs[5];  // 1/s and s as floats
  int n;
  void push(const float* ptr, int scale, int ch, int Hp, int Wp, const Tap* taps = nullptr) {
    xt[n] = taps; yt[n] = taps ? taps + Wp : nullptr;
    mf[n] = nullptr; f7[n] = nullptr;
    p[n] = ptr; s[n] = scale; c4[n] = ch / 4; Hl[n] = Hp / scale; Wl[n] = Wp / scale;
    rs[n] = 1.0f / (float)scale; fs[n] = (float)scale;
    n++;
  }
  void push_planes(const LowPlanes& pl, int scale, int Hp, int Wp, const Tap* taps = nullptr) {
    push(pl.flow, scale, 4, Hp, Wp, taps);
    mf[n - 1] = pl.mf; f7[n - 1] = pl.f7;
  }
};
// block 3 (s = 2) input gather fused with its conv0[0] (csrc/elementwise.cu)
int k_gather_s2_conv0a(vfi_handle* h, const Layer& Lb0, const uint8_t* imgs, const void* feats, const LowSet& lows,
                       const float* tsteps_dev, bf16* out, int B, int I1, int T, int Hp, int Wp, cudaStream_t s);
// tap tables of `n` levels: level k at out + k * (Wp + Hp), columns first
int k_build_taps(vfi_handle* h, Tap* out, const int* scales, int n, int Hp, int Wp, cudaStream_t s);
int k_gather(vfi_handle* h, const uint8_t* imgs, const void* feats, const LowSet& lows, const float* tsteps_dev,
             void* out, int B, int I1, int T, int Hp, int Wp, int s_int, bool bf, bool phase, cudaStream_t s);
// I1: index of pair 0's second frame in imgs / feats (B when the two frame sets are stored one after the other, 1 when
// the pairs are consecutive frames of one clip)
// flow_pre (nullable): [N,Hp,Wp,4] fp32 flow already summed over every level of `lows` but the last (full-resolution) one
int k_final_blend(vfi_handle* h, const uint8_t* imgs, const LowSet& lows, const float* flow_pre, const uint8_t* hud,
                  const float* tsteps_dev, uint8_t* out, float* out_f32, float* flow_out, int B, int I1, int T, int H, int W, int Hp, int Wp,
                  cudaStream_t s, int bgr = 0);
// finest block at scale 1.0: gather fused with conv0[0] (tcgen05), bf16 only
int k_gather_conv0a(vfi_handle* h, const Layer& Lb0, const uint8_t* imgs, const void* feats, const LowSet& lows,
                    const float* tsteps_dev, bf16* out, float* flow_sum, int B, int I1, int T, int Hp, int Wp,
                    cudaStream_t s);
int k_pack_fm(vfi_handle* h, const float* flow, const float* mask, float* fm, int N, int H, int W, int Hp, int Wp,
              cudaStream_t s);
int k_f32_to_act(vfi_handle* h, const float* src, void* dst, size_t n, bool bf, cudaStream_t s);
int k_act_to_f32(vfi_handle* h, const void* src, float* dst, size_t n, bool bf, cudaStream_t s);
int k_plain_to_phase(vfi_handle* h, const bf16* src, bf16* dst, int N, int H, int W, int C, cudaStream_t s);

// ---- conv_fp32.cu -------------------------------------------------------------------------
// in/out element type: float (bf=false) or bf16 (bf=true); accumulation always fp32.
int k_conv3_direct(vfi_handle* h, const Layer& L, const void* in, const void* residual, void* out, int N, int Hin,
                   int Win, bool bf, float slope, cudaStream_t s);
// mode 0: out [N,2Hi,2Wi,cout_pad] (act type); mode 1: pixel-shuffled fp32 low13 [N,4Hi,4Wi,16]
int k_deconv4_direct(vfi_handle* h, const Layer& L, const void* in, void* out, int N, int Hin, int Win, bool bf,
                     int mode, cudaStream_t s);

// ---- conv_tc.cu ---------------------------------------------------------------------------
int tc_build_layer(vfi_handle* h, Layer& L, const std::vector<float>& w_torch);  // packs wtc + step table
// conv (mode 0/1) : in plain (stride 1) or phase planes (stride 2) bf16, out bf16 plain or phase (out_phase)
// lastconv (deconv): in plain bf16, out = fp32 low [N,4Hin,4Win,low_c]: low_c = 16 (13 channels) or 8 (the final
// block: only flow delta + mask, 5 channels, are consumed downstream)
// planes (lastconv of blocks 0..3, bf16 mode): write the level as LowPlanes instead of fp32 texels
int k_conv_tc(vfi_handle* h, const Layer& L, const bf16* in, const bf16* residual, void* out, int N, int Hin,
              int Win, bool out_phase, float slope, cudaStream_t s, int low_c = 16, const LowPlanes* planes = nullptr);
// plain (no swizzle) 4-D tensor map over a bf16 NHWC tensor, box {bc, bw, bh, 1}, zero fill outside
// 128-byte-swizzled map, box {64 ch, bw, bh, 1}
int tc_make_tmap_sw128(vfi_handle* h, void* out_tmap, const bf16* base, int C, int W, int H, int N, int bw, int bh);
int tc_make_tmap_plain(vfi_handle* h, void* out_tmap, const bf16* base, int C, int W, int H, int N, int bc, int bw, int bh);
int k_unpack_planes(vfi_handle* h, const LowPlanes& pl, float* out16, size_t texels, cudaStream_t s);

// ---- conv_v2.cu ---------------------------------------------------------------------------
bool v2_supported(const Layer& L);
int v2_build_layer(vfi_handle* h, Layer& L, const std::vector<float>& w_torch);
// stride-1 3x3 conv (+ResConv epilogue), plain NHWC bf16 in/out
int k_conv_v2(vfi_handle* h, const Layer& L, const bf16* in, bf16* out, int N, int H, int W, float slope,
              cudaStream_t s);

int k_conv_v2_last(vfi_handle* h, const Layer& L, const bf16* in, float* low, int low_c, int N, int H, int W,
                   cudaStream_t s);

// ---- conv_s2.cu ---------------------------------------------------------------------------
bool s2_supported(const Layer& L);
int s2_build_layer(vfi_handle* h, Layer& L, const std::vector<float>& w_torch);
// stride-2 3x3 conv + LeakyReLU: phase-plane bf16 in, plain or phase-plane bf16 out
int k_conv_s2(vfi_handle* h, const Layer& L, const bf16* in, bf16* out, int N, int Hin, int Win, bool out_phase,
              float slope, cudaStream_t s);

// ---- reduce.cu is reached through the ABI functions; farneback.cu chains the flow statistics --------------------
int vfi_flow_stats_impl(vfi_handle* h, const float* flow, int n, int H, int W, double* out, cudaStream_t s);
