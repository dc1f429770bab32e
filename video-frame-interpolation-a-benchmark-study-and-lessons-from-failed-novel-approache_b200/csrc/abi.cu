// abi.cu -- the extern "C" boundary of libvfi_b200.so: handle lifetime, IFNet weight packing and the
// inference schedule (Head once per pair, 5 IFBlocks batched over pairs x timesteps, fused tail).
#include <math.h>

#include "common.cuh"

static const int BLOCK_C[5] = {192, 128, 96, 64, 32};

int vfi_ws_reserve(vfi_handle* h, size_t bytes) {
  if (bytes <= h->ws_bytes) return 0;
  if (h->ws) { cudaDeviceSynchronize(); cudaFree(h->ws); h->ws = nullptr; h->ws_bytes = 0; }
  size_t want = bytes + bytes / 8;
  if (cudaMalloc(&h->ws, want) != cudaSuccess) {
    cudaGetLastError();
    VFI_FAIL(h, VFI_E_NOMEM, "workspace cudaMalloc(%zu) failed", want);
  }
  h->ws_bytes = want;
  return 0;
}
int vfi_red_reserve(vfi_handle* h, size_t bytes) {
  if (bytes <= h->red_bytes) return 0;
  if (h->red) { cudaDeviceSynchronize(); cudaFree(h->red); h->red = nullptr; h->red_bytes = 0; }
  if (cudaMalloc(&h->red, bytes) != cudaSuccess) {
    cudaGetLastError();
    VFI_FAIL(h, VFI_E_NOMEM, "scratch cudaMalloc(%zu) failed", bytes);
  }
  h->red_bytes = bytes;
  return 0;
}

extern "C" __attribute__((visibility("default"))) int vfi_create(int device, vfi_handle** out) {
  if (!out) return VFI_E_ARG;
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) { cudaGetLastError(); return VFI_E_CUDA; }
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return VFI_E_CUDA;
  if (prop.major != 10) return VFI_E_CUDA;  // sm_100a only: no fallback path exists
  int prev = -1;
  cudaGetDevice(&prev);
  if (cudaSetDevice(device) != cudaSuccess) return VFI_E_CUDA;
  vfi_handle* h = new vfi_handle();
  h->device = device;
  h->sm_count = prop.multiProcessorCount;
  for (int i = 0; i < VFI_NUM_LAYERS; i++) h->layers[i] = Layer();
  const bool ok = cudaEventCreateWithFlags(&h->scratch_ev, cudaEventDisableTiming) == cudaSuccess &&
                  cudaMalloc(&h->tdev, 64 * sizeof(float)) == cudaSuccess && cudaMalloc(&h->hud_cnt, 16) == cudaSuccess &&
                  cudaMemset(h->hud_cnt, 0, 16) == cudaSuccess;
  if (prev >= 0 && prev != device) cudaSetDevice(prev);  // leave the caller's current device as it was
  if (!ok) { cudaGetLastError(); delete h; return VFI_E_CUDA; }
  *out = h;
  return VFI_OK;
}

static void free_layers(vfi_handle* h) {
  for (int i = 0; i < VFI_NUM_LAYERS; i++) {
    Layer& L = h->layers[i];
    cudaFree(L.w32); cudaFree(L.bias); cudaFree(L.beta); cudaFree(L.wtc); cudaFree(L.bias_tc); cudaFree(L.wv2); cudaFree(L.bias_v2); cudaFree(L.ws2);
    L.w32 = L.bias = L.beta = L.bias_tc = L.bias_v2 = nullptr;
    L.wtc = nullptr;
    L.wv2 = nullptr;
    L.ws2 = nullptr;
  }
}

extern "C" __attribute__((visibility("default"))) void vfi_destroy(vfi_handle* h) {
  if (!h) return;
  int prev = -1;
  cudaGetDevice(&prev);
  cudaSetDevice(h->device);
  cudaDeviceSynchronize();
  free_layers(h);
  cudaFree(h->ws);
  cudaFree(h->red);
  cudaFree(h->tdev);
  cudaFree(h->hud_cnt);
  if (h->scratch_ev) cudaEventDestroy(h->scratch_ev);
  if (prev >= 0 && prev != h->device) cudaSetDevice(prev);
  delete h;
}

void prof_begin(vfi_handle* h, int cat, cudaStream_t s) {
  cudaEvent_t a, b;
  cudaEventCreate(&a);
  cudaEventCreate(&b);
  h->prof_ev.push_back(a);
  h->prof_ev.push_back(b);
  h->prof_cat.push_back(cat);
  cudaEventRecord(a, s);
}
void prof_end(vfi_handle* h, cudaStream_t s) { cudaEventRecord(h->prof_ev.back(), s); }

extern "C" __attribute__((visibility("default"))) int vfi_profile_enable(vfi_handle* h, int on) {
  if (!h) return VFI_E_ARG;
  h->profile = on != 0;
  return VFI_OK;
}
// Synchronises, sums the elapsed ms and launch counts per kernel class (ProfCat order) and clears the log.
extern "C" __attribute__((visibility("default"))) int vfi_profile_read(vfi_handle* h, double* ms, long long* counts, int ncat) {
  if (!h || !ms || !counts) return VFI_E_ARG;
  for (int i = 0; i < ncat; i++) { ms[i] = 0; counts[i] = 0; }
  VFI_CUDA(h, cudaDeviceSynchronize());
  for (size_t i = 0; i < h->prof_cat.size(); i++) {
    float t = 0.f;
    cudaEventElapsedTime(&t, h->prof_ev[2 * i], h->prof_ev[2 * i + 1]);
    int c = h->prof_cat[i];
    if (c < ncat) { ms[c] += t; counts[c]++; }
    cudaEventDestroy(h->prof_ev[2 * i]);
    cudaEventDestroy(h->prof_ev[2 * i + 1]);
  }
  h->prof_ev.clear();
  h->prof_cat.clear();
  return VFI_OK;
}

extern "C" __attribute__((visibility("default"))) int vfi_set_option(vfi_handle* h, const char* key, int value) {
  if (!h || !key) return VFI_E_ARG;
  if (!strcmp(key, "conv_v2")) { h->opt_conv_v2 = value; return VFI_OK; }
  if (!strcmp(key, "v2_ablate")) { h->opt_v2_ablate = value; return VFI_OK; }
  if (!strcmp(key, "fuse_gather")) { h->opt_fuse_gather = value; return VFI_OK; }
  if (!strcmp(key, "v2_res_global")) { h->opt_v2_res_global = value; return VFI_OK; }
  if (!strcmp(key, "v2_pm")) { h->opt_v2_pm = value; return VFI_OK; }
  if (!strcmp(key, "v2_tma")) { h->opt_v2_tma = value; return VFI_OK; }
  if (!strcmp(key, "l2_chunk_mb")) { h->opt_l2_chunk_mb = value; return VFI_OK; }
  if (!strcmp(key, "fuse_gather3")) { h->opt_fuse_gather3 = value; return VFI_OK; }
  if (!strcmp(key, "gc_persistent")) { h->opt_gc_persistent = value; return VFI_OK; }
  if (!strcmp(key, "t_fast")) { h->opt_t_fast = value; return VFI_OK; }
  if (!strcmp(key, "gc_ablate")) { h->opt_gc_ablate = value; return VFI_OK; }
  if (!strcmp(key, "v2_trace")) { h->opt_v2_trace = value; return VFI_OK; }
  if (!strcmp(key, "ws_guard")) { h->opt_ws_guard = value; return VFI_OK; }
  VFI_FAIL(h, VFI_E_ARG, "unknown option %s", key);
}

// Orders `stream` into the handle's call sequence without launching anything: waits (on the device) for the handle's previous
// call when that ran on another stream, and marks `stream` as the latest user.  Bracket a CUDA-graph replay of captured
// calls with two fences (captured calls record no cross-stream bookkeeping themselves).
extern "C" __attribute__((visibility("default"))) int vfi_fence(vfi_handle* h, void* stream) {
  if (!h) return VFI_E_ARG;
  VFI_ENTER(h, stream);
  return VFI_OK;
}
extern "C" __attribute__((visibility("default"))) const char* vfi_last_error(vfi_handle* h) { return h ? h->err.c_str() : "null handle"; }
extern "C" __attribute__((visibility("default"))) long long vfi_launch_count(vfi_handle* h) { return h ? h->launches : 0; }
extern "C" __attribute__((visibility("default"))) size_t vfi_device_bytes(vfi_handle* h) {
  return h ? h->ws_bytes + h->red_bytes + h->weight_bytes + 64 * sizeof(float) : 0;
}

static void layer_shape(int id, Layer& L) {
  L = Layer();
  L.act = true;
  L.has_beta = false;
  L.stride = 1;
  L.kind = LK_CONV3;
  if (id < 4) {
    if (id == 0) { L.cin = 3; L.cout = 16; L.stride = 2; L.cin_pad = 4; }
    else if (id < 3) { L.cin = 16; L.cout = 16; L.cin_pad = 16; }
    else { L.kind = LK_DECONV4; L.cin = 16; L.cout = 4; L.cin_pad = 16; L.act = false; L.stride = 2; }
  } else {
    int b = (id - 4) / 11, j = (id - 4) % 11, c = BLOCK_C[b];
    if (j == 0) { L.cin = b == 0 ? 15 : 28; L.cin_pad = b == 0 ? 16 : 32; L.cout = c / 2; L.stride = 2; }
    else if (j == 1) { L.cin = L.cin_pad = c / 2; L.cout = c; L.stride = 2; }
    else if (j < 10) { L.cin = L.cin_pad = c; L.cout = c; L.has_beta = true; }
    else { L.kind = LK_DECONV4; L.cin = L.cin_pad = c; L.cout = 52; L.act = false; L.stride = 2; }
  }
  L.cout_pad = L.cout;
}

extern "C" __attribute__((visibility("default"))) int vfi_ifnet_load_weights(vfi_handle* h, const float* params, size_t n_params, const float* lrelu_slope) {
  if (!h) return VFI_E_ARG;
  if (!params || n_params != (size_t)VFI_IFNET_PARAMS)
    VFI_FAIL(h, VFI_E_ARG, "expected %d parameters, got %zu", VFI_IFNET_PARAMS, n_params);
  VFI_ENTER(h, 0);
  free_layers(h);
  h->weights_loaded = false;
  size_t free0 = 0, free1 = 0, total = 0;
  cudaMemGetInfo(&free0, &total);
  h->slope = lrelu_slope ? *lrelu_slope : 0.2f;
  size_t off = 0;
  for (int id = 0; id < VFI_NUM_LAYERS; id++) {
    Layer& L = h->layers[id];
    layer_shape(id, L);
    const int kk = L.kind == LK_CONV3 ? 9 : 16;
    size_t nw = (size_t)L.cin * L.cout * kk;
    if (off + nw + L.cout + (L.has_beta ? L.cout : 0) > n_params) VFI_FAIL(h, VFI_E_ARG, "parameter blob too short");
    std::vector<float> w(params + off, params + off + nw);
    off += nw;
    std::vector<float> bias(params + off, params + off + L.cout);
    off += L.cout;
    std::vector<float> beta;
    if (L.has_beta) { beta.assign(params + off, params + off + L.cout); off += L.cout; }
    // fp32 layout [tap][cin_pad][cout_pad]
    std::vector<float> w32((size_t)kk * L.cin_pad * L.cout_pad, 0.f);
    for (int co = 0; co < L.cout; co++)
      for (int ci = 0; ci < L.cin; ci++)
        for (int t = 0; t < kk; t++) {
          float v = L.kind == LK_CONV3 ? w[((size_t)co * L.cin + ci) * 9 + t] : w[((size_t)ci * L.cout + co) * 16 + t];
          w32[((size_t)t * L.cin_pad + ci) * L.cout_pad + co] = v;
        }
    if (id == 0) {
      for (int t = 0; t < 9; t++)
        for (int ci = 0; ci < 3; ci++)
          for (int co = 0; co < 16; co++) h->head0_w[(t * 3 + ci) * 16 + co] = w[((size_t)co * 3 + ci) * 9 + t] / 255.0f;
      for (int co = 0; co < 16; co++) h->head0_b[co] = bias[co];
    }
    VFI_CUDA(h, cudaMalloc(&L.w32, w32.size() * 4));
    VFI_CUDA(h, cudaMemcpy(L.w32, w32.data(), w32.size() * 4, cudaMemcpyHostToDevice));
    std::vector<float> bpad(64 > L.cout_pad ? 64 : L.cout_pad, 0.f);
    for (int i = 0; i < L.cout; i++) bpad[i] = bias[i];
    VFI_CUDA(h, cudaMalloc(&L.bias, bpad.size() * 4));
    VFI_CUDA(h, cudaMemcpy(L.bias, bpad.data(), bpad.size() * 4, cudaMemcpyHostToDevice));
    if (L.has_beta) {
      VFI_CUDA(h, cudaMalloc(&L.beta, beta.size() * 4));
      VFI_CUDA(h, cudaMemcpy(L.beta, beta.data(), beta.size() * 4, cudaMemcpyHostToDevice));
    }
    const bool tc = (id == 1 || id == 2 || id >= 4);
    if (tc) {
      if (int rc = tc_build_layer(h, L, w)) return rc;
      if (int rc = s2_build_layer(h, L, w)) return rc;
    }
    if (tc || id == 3) {
      if (id != 58 && L.kind == LK_DECONV4 && L.cout == 52) { /* only the final block's lastconv has a conv_v2 form */ }
      else if (int rc = v2_build_layer(h, L, w)) return rc;
    }
  }
  if (off != n_params) VFI_FAIL(h, VFI_E_ARG, "parameter blob length mismatch (%zu consumed)", off);
  cudaMemGetInfo(&free1, &total);
  h->weight_bytes = free0 > free1 ? free0 - free1 : 0;  // what the packed weight images actually took from the device
  h->weights_loaded = true;
  return VFI_OK;
}

// ---------------------------------------------------------------------------------------------
// Debug check behind vfi_set_option("ws_guard", 1) (tests/test_gpu_bounds.py): the workspace is filled with WS_GUARD_BYTE before
// the schedule runs, every carved buffer is followed by a gap no kernel may write, and this kernel counts the gap bytes that
// changed (bad[0]) and keeps the lowest gap index among them (bad[1]).
__global__ void ws_guard_check_kernel(const uint8_t* __restrict__ ws, const unsigned long long* __restrict__ gaps, int ngaps,
                                      unsigned long long* __restrict__ bad) {
  const int g = blockIdx.x;
  if (g >= ngaps) return;
  const uint8_t* p = ws + gaps[2 * g];
  const unsigned long long len = gaps[2 * g + 1];
  unsigned long long mine = 0;
  for (unsigned long long i = threadIdx.x; i < len; i += blockDim.x) mine += p[i] != WS_GUARD_BYTE;
  if (mine) { atomicAdd(&bad[0], mine); atomicMin(&bad[1], (unsigned long long)g); }
}
int vfi_ws_guard_check(vfi_handle* h, const std::vector<unsigned long long>& gaps, unsigned long long* dev /* 2 + gaps */,
                          cudaStream_t s) {
  const int ngaps = (int)(gaps.size() / 2);
  const unsigned long long init[2] = {0ull, ~0ull};
  VFI_CUDA(h, cudaMemcpyAsync(dev, init, sizeof(init), cudaMemcpyHostToDevice, s));
  VFI_CUDA(h, cudaMemcpyAsync(dev + 2, gaps.data(), gaps.size() * sizeof(unsigned long long), cudaMemcpyHostToDevice, s));
  ws_guard_check_kernel<<<ngaps, 256, 0, s>>>((const uint8_t*)h->ws, dev + 2, ngaps, dev);
  unsigned long long bad[2];
  VFI_CUDA(h, cudaMemcpyAsync(bad, dev, sizeof(bad), cudaMemcpyDeviceToHost, s));
  VFI_CUDA(h, cudaStreamSynchronize(s));
  if (bad[0]) VFI_FAIL(h, VFI_E_CUDA, "workspace guard: %llu byte(s) written past the end of carved buffer %llu (of %d)", bad[0], bad[1], ngaps);
  return VFI_OK;
}

static int run_conv(vfi_handle* h, const Layer& L, const void* in, const void* res, void* out, int N, int Hin,
                    int Win, bool bf, bool out_phase, cudaStream_t s) {
  if (bf && h->opt_conv_v2 && L.wv2 && !out_phase && (res == nullptr || res == in))
    return k_conv_v2(h, L, (const bf16*)in, (bf16*)out, N, Hin, Win, h->slope, s);
  if (bf && h->opt_conv_v2 && L.ws2 && res == nullptr)
    return k_conv_s2(h, L, (const bf16*)in, (bf16*)out, N, Hin, Win, out_phase, h->slope, s);
  if (bf) return k_conv_tc(h, L, (const bf16*)in, (const bf16*)res, out, N, Hin, Win, out_phase, h->slope, s);
  return k_conv3_direct(h, L, in, res, out, N, Hin, Win, false, h->slope, s);
}

extern "C" __attribute__((visibility("default"))) int vfi_ifnet_infer(vfi_handle* h, const uint8_t* img0, const uint8_t* img1, int B, int H, int W,
                               const float* timesteps, int T, float scale, int precision, const uint8_t* hud_mask,
                               uint8_t* out, float* out_f32, float* flow_out, void* stream) {
  return vfi_ifnet_infer_ex(h, img0, img1, B, H, W, timesteps, T, scale, precision, 0, hud_mask, out, out_f32, flow_out, stream);
}
extern "C" __attribute__((visibility("default"))) int vfi_ifnet_infer_ex(vfi_handle* h, const uint8_t* img0, const uint8_t* img1, int B, int H, int W,
                                  const float* timesteps, int T, float scale, int precision, int flags, const uint8_t* hud_mask,
                                  uint8_t* out, float* out_f32, float* flow_out, void* stream) {
  if (!h) return VFI_E_ARG;
  if (flags & ~(VFI_FLAG_BGR | VFI_FLAG_PAD_REFLECT_EDGE)) VFI_FAIL(h, VFI_E_ARG, "unknown flags 0x%x", flags);
  const int bgr = (flags & VFI_FLAG_BGR) ? 1 : 0, edge = (flags & VFI_FLAG_PAD_REFLECT_EDGE) ? 1 : 0;
  if (!h->weights_loaded) VFI_FAIL(h, VFI_E_STATE, "IFNet weights not loaded");
  if (!img0 || !img1 || !out || !timesteps || B < 1 || T < 1 || T > 64 || H < 2 || W < 2)
    VFI_FAIL(h, VFI_E_ARG, "bad argument");
  if (precision != VFI_PRECISION_FP32 && precision != VFI_PRECISION_BF16) VFI_FAIL(h, VFI_E_ARG, "bad precision");
  int inv;
  if (scale == 1.0f) inv = 1; else if (scale == 0.5f) inv = 2; else if (scale == 0.25f) inv = 4;
  else VFI_FAIL(h, VFI_E_ARG, "scale must be 1.0, 0.5 or 0.25");
  VFI_ENTER(h, stream);
  cudaStream_t s = (cudaStream_t)stream;
  const bool bf = precision == VFI_PRECISION_BF16;
  const size_t esz = bf ? 2 : 4;
  // run_experiments.py:373-375 pads to multiples of 128 at every scale; the wrapper (base.py:203-225) to max(32, 64/scale)
  const int mult = edge ? std::max(128, 64 * inv) : (64 * inv > 32 ? 64 * inv : 32);
  const int Hp = (H + mult - 1) / mult * mult, Wp = (W + mult - 1) / mult * mult;
  if (Hp - H >= H + edge || Wp - W >= W + edge) VFI_FAIL(h, VFI_E_ARG, "frame too small for reflect padding to %d", mult);
  // consecutive pairs of one clip (img1 is img0 shifted by one frame): frame b+1 is both pair b's second and pair
  // b+1's first frame -- pad it and run the Head on it once
  const bool clipmode = B > 1 && img1 == img0 + (size_t)H * W * 3;
  const int N = B * T, NI = clipmode ? B + 1 : 2 * B, I1 = clipmode ? 1 : B;
  // the gather / blend kernels index pixels of the whole image set with 32-bit integers
  if ((size_t)std::max(NI, N) * Hp * Wp >= ((size_t)1 << 31))
    VFI_FAIL(h, VFI_E_ARG, "batch too large: %d images of %dx%d exceed 2^31 pixels, split the call", std::max(NI, N), Hp, Wp);
  int sl[5];
  for (int i = 0; i < 5; i++) sl[i] = (16 >> i) * inv;

  // ---- workspace ----
  size_t xin_e = 0, a0_e = 0, f_e = 0;
  for (int i = 0; i < 5; i++) {
    size_t hs = Hp / sl[i], wsz = Wp / sl[i];
    int c = BLOCK_C[i];
    xin_e = std::max(xin_e, (size_t)N * hs * wsz * (i == 0 ? 16 : 32));
    a0_e = std::max(a0_e, (size_t)N * (hs / 2) * (wsz / 2) * (c / 2));
    f_e = std::max(f_e, (size_t)N * (hs / 4) * (wsz / 4) * c);
  }
  // the final block's lastconv output keeps only flow delta + mask (8-float texels) on the tcgen05 path
  // ... and when the finest block's gather is the fused kernel (which leaves the coarser blocks' flow summed per pixel for
  // the blend), as fp16: 16-byte texels
  const Layer* Lf = &h->layers[4 + 11 * 4];
  const bool fused = bf && inv == 1 && h->opt_fuse_gather && h->opt_conv_v2 && Lf[0].ws2 && Lf[0].cout == 16;
  const int last_c = !bf ? 16 : (fused && Lf[10].wv2) ? 4 : 8;
  const size_t px = (size_t)Hp * Wp;
  Carve cv{nullptr, 0};
  std::vector<unsigned long long> guard_gaps;
  if (h->opt_ws_guard) { cv.guard = WS_GUARD_BYTES; cv.gaps = &guard_gaps; }
  for (int pass = 0; pass < 2; pass++) {
    cv.off = 0;
    cv.base = pass ? (uint8_t*)h->ws : nullptr;
    if (pass == 0) { /* sizing pass */ }
    if (cv.guard) cv.take<uint8_t>(0);                    // a gap before the first buffer too
    uint8_t* imgs = cv.take<uint8_t>(NI * px * 4);
    uint8_t* feats = cv.take<uint8_t>(NI * px * 4 * esz);
    uint8_t* ht0 = cv.take<uint8_t>(NI * (px / 4) * 16 * esz);
    uint8_t* ht1 = cv.take<uint8_t>(NI * (px / 4) * 16 * esz);
    float* tdev = h->tdev;
    Tap* taps = cv.take<Tap>((size_t)5 * (Hp + Wp));
    float* flow3 = cv.take<float>((size_t)N * px * 4);  // blocks 0-3 flow per pixel (written by the fused finest gather)
    // coarse level i = block i's lastconv output: LowPlanes (blocks 0..3, bf16 mode) or fp32 texels
    float* low[5];
    LowPlanes lowp[5];
    for (int i = 0; i < 5; i++) {
      const size_t texels = (size_t)N * (Hp / sl[i]) * (Wp / sl[i]);
      if (bf && i < 4) {
        lowp[i].flow = cv.take<float>(texels * 4);
        lowp[i].mf = cv.take<bf16>(texels * 8);
        lowp[i].f7 = cv.take<bf16>(texels);
        low[i] = nullptr;
      } else {
        low[i] = cv.take<float>(texels * (i == 4 ? last_c : 16));
      }
    }
    uint8_t* xin = cv.take<uint8_t>(xin_e * esz);
    uint8_t* a0 = cv.take<uint8_t>(a0_e * esz);
    uint8_t* fa = cv.take<uint8_t>(f_e * esz);
    uint8_t* fb = cv.take<uint8_t>(f_e * esz);
    const size_t guarded_bytes = cv.off;
    unsigned long long* guard_dev = nullptr;
    if (cv.guard) guard_dev = ws_guard_scratch(cv);        // the check's own counters + gap list, outside the guarded range
    if (pass == 0) {
      if (int rc = vfi_ws_reserve(h, cv.off + 4096)) return rc;
      continue;
    }
    int rc;
    if (cv.guard) {
      if (guard_gaps.size() > 2 * WS_GUARD_MAX_GAPS) VFI_FAIL(h, VFI_E_ARG, "ws_guard: too many carved buffers");
      VFI_CUDA(h, cudaMemsetAsync(h->ws, WS_GUARD_BYTE, guarded_bytes, s));
    }
    if (T != h->t_n || memcmp(h->t_host, timesteps, T * sizeof(float))) {  // steady state (same timesteps every call): no copy
      memcpy(h->t_host, timesteps, T * sizeof(float));
      h->t_n = 0;  // invalid until the copy is enqueued
      VFI_CUDA(h, cudaMemcpyAsync(tdev, timesteps, T * sizeof(float), cudaMemcpyHostToDevice, s));
      h->t_n = T;
    }
    if ((rc = k_build_taps(h, taps, sl, 5, Hp, Wp, s))) return rc;
    // to_tensor + reflect pad, both frames of every pair
    if ((rc = k_pad_rgbx(h, img0, imgs, B, H, W, Hp, Wp, s, edge, bgr))) return rc;
    if (clipmode) rc = k_pad_rgbx(h, img1 + (size_t)(B - 1) * H * W * 3, imgs + (size_t)B * px * 4, 1, H, W, Hp, Wp, s, edge, bgr);
    else rc = k_pad_rgbx(h, img1, imgs + (size_t)B * px * 4, B, H, W, Hp, Wp, s, edge, bgr);
    if (rc) return rc;
    // Head (encode) once per image, shared by all T timesteps
    if ((rc = k_head_cnn0(h, h->layers[0], imgs, ht0, NI, Hp, Wp, bf, s))) return rc;
    if ((rc = run_conv(h, h->layers[1], ht0, nullptr, ht1, NI, Hp / 2, Wp / 2, bf, false, s))) return rc;
    if ((rc = run_conv(h, h->layers[2], ht1, nullptr, ht0, NI, Hp / 2, Wp / 2, bf, false, s))) return rc;
    if (bf && h->opt_conv_v2) rc = k_conv_v2(h, h->layers[3], (const bf16*)ht0, (bf16*)feats, NI, Hp / 2, Wp / 2, h->slope, s);
    else rc = k_deconv4_direct(h, h->layers[3], ht0, feats, NI, Hp / 2, Wp / 2, bf, 0, s);
    if (rc) return rc;
    LowSet lows;
    memset(&lows, 0, sizeof(lows));
    const float* flow_pre = nullptr;
    for (int i = 0; i < 5; i++) {
      const int sc = sl[i], hs = Hp / sc, wsz = Wp / sc;
      const Layer* Lb = &h->layers[4 + 11 * i];
      if (fused && i == 4) {
        if ((rc = k_gather_conv0a(h, Lb[0], imgs, feats, lows, tdev, (bf16*)a0, flow3, B, I1, T, Hp, Wp, s))) return rc;
        flow_pre = flow3;
      } else if (fused && h->opt_fuse_gather3 && i == 3 && sc == 2 && Lb[0].ws2 && Lb[0].cout == 32 && Hp % 64 == 0 && Wp % 32 == 0) {
        if ((rc = k_gather_s2_conv0a(h, Lb[0], imgs, feats, lows, tdev, (bf16*)a0, B, I1, T, Hp, Wp, s))) return rc;
      } else {
        if ((rc = k_gather(h, imgs, feats, lows, tdev, xin, B, I1, T, Hp, Wp, sc, bf, bf, s))) return rc;
        if ((rc = run_conv(h, Lb[0], xin, nullptr, a0, N, hs, wsz, bf, bf, s))) return rc;
      }
      if ((rc = run_conv(h, Lb[1], a0, nullptr, fa, N, hs / 2, wsz / 2, bf, false, s))) return rc;
      uint8_t *cur = fa, *nxt = fb;
      {
        // ResConv chain, depth-first over groups of items: a group's activations (input + output of one layer) stay
        // under opt_l2_chunk_mb, so layer j+1 reads from L2 what layer j just wrote instead of going to HBM
        const size_t item_bytes = (size_t)(hs / 4) * (wsz / 4) * BLOCK_C[i] * esz;
        int grp = N;
        if (bf && h->opt_l2_chunk_mb > 0) grp = (int)std::max<size_t>(1, std::min<size_t>(N, ((size_t)h->opt_l2_chunk_mb << 20) / (2 * item_bytes)));
        for (int n0 = 0; n0 < N; n0 += grp) {
          const int nn = std::min(grp, N - n0);
          uint8_t *c0 = cur + (size_t)n0 * item_bytes, *c1 = nxt + (size_t)n0 * item_bytes;
          for (int j = 0; j < 8; j++) {
            if ((rc = run_conv(h, Lb[2 + j], c0, c0, c1, nn, hs / 4, wsz / 4, bf, false, s))) return rc;
            std::swap(c0, c1);
          }
        }
        // eight swaps: the chain's result is back in `cur`
      }
      const int lc = i == 4 ? last_c : 16;
      if (bf && i == 4 && h->opt_conv_v2 && Lb[10].wv2) rc = k_conv_v2_last(h, Lb[10], (const bf16*)cur, low[i], lc, N, hs / 4, wsz / 4, s);
      else if (bf && i < 4) rc = k_conv_tc(h, Lb[10], (const bf16*)cur, nullptr, nullptr, N, hs / 4, wsz / 4, false, h->slope, s, 16, &lowp[i]);
      else if (bf) rc = k_conv_tc(h, Lb[10], (const bf16*)cur, nullptr, low[i], N, hs / 4, wsz / 4, false, h->slope, s, lc);
      else rc = k_deconv4_direct(h, Lb[10], cur, low[i], N, hs / 4, wsz / 4, false, 1, s);
      if (rc) return rc;
      if (bf && i < 4) lows.push_planes(lowp[i], sc, Hp, Wp, taps + (size_t)i * (Hp + Wp));
      else lows.push(low[i], sc, lc, Hp, Wp, taps + (size_t)i * (Hp + Wp));
    }
    if ((rc = k_final_blend(h, imgs, lows, flow_pre, hud_mask, tdev, out, out_f32, flow_out, B, I1, T, H, W, Hp, Wp, s, bgr))) return rc;
    if (cv.guard) {
      // value 2 = the check's self-test: one byte of the fourth gap is overwritten on purpose and the call must fail
      if (h->opt_ws_guard == 2) VFI_CUDA(h, cudaMemsetAsync((uint8_t*)h->ws + guard_gaps[2 * 3], 0, 1, s));
      if ((rc = vfi_ws_guard_check(h, guard_gaps, guard_dev, s))) return rc;
    }
  }
  return VFI_OK;
}

extern "C" __attribute__((visibility("default"))) int vfi_warp_blend(vfi_handle* h, const uint8_t* img0, const uint8_t* img1, const float* flow,
                              const float* mask, int N, int H, int W, uint8_t* out, float* out_f32, void* stream) {
  if (!h) return VFI_E_ARG;
  if (!img0 || !img1 || !flow || !mask || !out || N < 1 || H < 1 || W < 1) VFI_FAIL(h, VFI_E_ARG, "bad argument");
  VFI_ENTER(h, stream);
  cudaStream_t s = (cudaStream_t)stream;
  const size_t px = (size_t)H * W;
  Carve cv{nullptr, 0};
  for (int pass = 0; pass < 2; pass++) {
    cv.off = 0;
    cv.base = pass ? (uint8_t*)h->ws : nullptr;
    uint8_t* imgs = cv.take<uint8_t>(2 * (size_t)N * px * 4);
    float* fm = cv.take<float>((size_t)N * px * 8);
    float* tdev = cv.take<float>(64);
    if (pass == 0) { if (int rc = vfi_ws_reserve(h, cv.off + 4096)) return rc; continue; }
    int rc;
    VFI_CUDA(h, cudaMemsetAsync(tdev, 0, 64 * sizeof(float), s));
    if ((rc = k_pad_rgbx(h, img0, imgs, N, H, W, H, W, s))) return rc;
    if ((rc = k_pad_rgbx(h, img1, imgs + (size_t)N * px * 4, N, H, W, H, W, s))) return rc;
    if ((rc = k_pack_fm(h, flow, mask, fm, N, H, W, H, W, s))) return rc;
    LowSet lows;
    memset(&lows, 0, sizeof(lows));
    lows.push(fm, 1, 8, H, W);
    if ((rc = k_final_blend(h, imgs, lows, nullptr, nullptr, tdev, out, out_f32, nullptr, N, N, 1, H, W, H, W, s))) return rc;
  }
  return VFI_OK;
}

// ---------------------------------------------------------------------------------------------
extern "C" __attribute__((visibility("default"))) int vfi_debug_conv(vfi_handle* h, int layer, int precision, const float* x, int N, int Hin, int Win,
                              const float* residual, float* y, void* stream) {
  if (!h) return VFI_E_ARG;
  if (!h->weights_loaded) VFI_FAIL(h, VFI_E_STATE, "IFNet weights not loaded");
  if (layer < 1 || layer >= VFI_NUM_LAYERS || !x || !y) VFI_FAIL(h, VFI_E_ARG, "bad layer/argument");
  VFI_ENTER(h, stream);
  cudaStream_t s = (cudaStream_t)stream;
  const Layer& L = h->layers[layer];
  const bool bf = precision == VFI_PRECISION_BF16;
  const bool s2 = L.kind == LK_CONV3 && L.stride == 2;
  const int Ho = L.kind == LK_DECONV4 ? 2 * Hin : (s2 ? Hin / 2 : Hin);
  const int Wo = L.kind == LK_DECONV4 ? 2 * Win : (s2 ? Win / 2 : Win);
  const size_t nin = (size_t)N * Hin * Win * L.cin_pad;
  const size_t nout = (size_t)N * Ho * Wo * L.cout_pad;
  if (!bf) {
    if (L.kind == LK_CONV3)
      return k_conv3_direct(h, L, x, L.has_beta && !residual ? x : residual, y, N, Hin, Win, false, h->slope, s);
    return k_deconv4_direct(h, L, x, y, N, Hin, Win, false, layer == 3 ? 0 : 1, s);
  }
  Carve cv{nullptr, 0};
  for (int pass = 0; pass < 2; pass++) {
    cv.off = 0;
    cv.base = pass ? (uint8_t*)h->ws : nullptr;
    bf16* xb = cv.take<bf16>(nin);
    bf16* xp = cv.take<bf16>(nin);
    bf16* rb = cv.take<bf16>(nout);
    bf16* yb = cv.take<bf16>(nout);
    const size_t dtex = L.kind == LK_DECONV4 ? (size_t)N * 16 * Hin * Win : 0;
    bf16* dmf = cv.take<bf16>(dtex * 8);
    bf16* df7 = cv.take<bf16>(dtex);
    if (pass == 0) { if (int rc = vfi_ws_reserve(h, cv.off + 4096)) return rc; continue; }
    int rc;
    if ((rc = k_f32_to_act(h, x, xb, nin, true, s))) return rc;
    if (layer == 3) {
      if (h->opt_conv_v2) rc = k_conv_v2(h, L, xb, yb, N, Hin, Win, h->slope, s);
      else rc = k_deconv4_direct(h, L, xb, yb, N, Hin, Win, true, 0, s);
      if (rc) return rc;
      return k_act_to_f32(h, yb, y, nout, true, s);
    }
    const bf16* in = xb;
    if (s2) { if ((rc = k_plain_to_phase(h, xb, xp, N, Hin, Win, L.cin_pad, s))) return rc; in = xp; }
    if (L.kind == LK_DECONV4 && layer == 58 && h->opt_conv_v2 && L.wv2)  // channels 0-4 only; 5-15 stay as the caller set them
      return k_conv_v2_last(h, L, in, y, 16, N, Hin, Win, s);
    if (L.kind == LK_DECONV4) {  // blocks 0..3: through the production LowPlanes epilogue, then unpacked to [.,16] fp32
      const size_t texels = (size_t)N * 4 * Hin * 4 * Win;
      LowPlanes pl;
      pl.flow = reinterpret_cast<float*>(yb);  // nout = texels * 52 / 4 bf16 >= 16 B per texel
      pl.mf = dmf; pl.f7 = df7;
      if ((rc = k_conv_tc(h, L, in, nullptr, nullptr, N, Hin, Win, false, h->slope, s, 16, &pl))) return rc;
      return k_unpack_planes(h, pl, y, texels, s);
    }
    const bf16* resp = nullptr;
    if (L.has_beta) {  // ResConv: residual defaults to the layer input (x), as in the net
      resp = xb;
      if (residual) { if ((rc = k_f32_to_act(h, residual, rb, nout, true, s))) return rc; resp = rb; }
    }
    if (h->opt_conv_v2 && L.wv2 && !s2 && (resp == nullptr || resp == xb)) {
      if ((rc = k_conv_v2(h, L, xb, yb, N, Hin, Win, h->slope, s))) return rc;
    } else if (h->opt_conv_v2 && L.ws2 && s2) {
      if ((rc = k_conv_s2(h, L, in, yb, N, Hin, Win, false, h->slope, s))) return rc;
    } else if ((rc = k_conv_tc(h, L, in, resp, yb, N, Hin, Win, false, h->slope, s))) return rc;
    return k_act_to_f32(h, yb, y, nout, true, s);
  }
  return VFI_OK;
}

extern "C" __attribute__((visibility("default"))) int vfi_route_decide(double ssim_score, double motion_mean, double motion_max, double motion_std,
                                double laplacian_var, double scene_thr, double particle_thr, double motion_lo,
                                double motion_hi, int* route, double* confidence, int* has_particles,
                                double* particle_score) {
  if (!route || !confidence) return VFI_E_ARG;
  int hp = 0;
  double ps = 0.0;
  if (ssim_score < scene_thr) {  // adaptive_pipeline.py:242-254
    *route = VFI_ROUTE_SCENE_CHANGE;
    *confidence = 1.0 - ssim_score;
  } else {
    double fs = motion_std / 20.0; if (fs > 1.0) fs = 1.0;           // :178
    double qs = laplacian_var / 500.0; if (qs > 1.0) qs = 1.0;       // :186
    ps = sqrt(fs * qs);                                              // :189
    hp = ps > particle_thr;
    if (hp || motion_max > motion_hi) {                              // :266-268
      *route = VFI_ROUTE_VFIMAMBA;
      double c = ps + motion_max / 50.0;
      *confidence = c < 1.0 ? c : 1.0;
    } else if (motion_mean < motion_lo) {                            // :269-271
      *route = VFI_ROUTE_RIFE;
      *confidence = 1.0 - motion_mean / motion_lo;
    } else {                                                         // :272-275
      *route = VFI_ROUTE_RIFE;
      *confidence = 0.7;
    }
  }
  if (has_particles) *has_particles = hp;
  if (particle_score) *particle_score = ps;
  return VFI_OK;
}
