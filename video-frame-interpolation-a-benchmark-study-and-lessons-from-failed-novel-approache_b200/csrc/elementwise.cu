// elementwise.cu -- the HBM-bound kernels of the IFNet pass: reflect-pad, Head cnn0, block-input
// gather (backward warp + bilinear /s + channel concat fused), bilinear xs upsample + flow
// accumulate, and the final fused backward-warp + sigmoid-blend + u8 store.
// Reference semantics: SURVEY.md Appendix A.2 (IFNet.forward / IFBlock.forward / warp()).
#include "common.cuh"
#include "umma.cuh"
#include <cuda_fp16.h>
#include <type_traits>

// ------------------------------------------------------------------------------------------
// small typed load/store helpers
template <typename T> struct Act;
template <> struct Act<float> {
  static __device__ __forceinline__ void load4(const float* p, float* v) {
    float4 a = *reinterpret_cast<const float4*>(p);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
  }
  static __device__ __forceinline__ void store8(float* p, const float* v) {
    reinterpret_cast<float4*>(p)[0] = make_float4(v[0], v[1], v[2], v[3]);
    reinterpret_cast<float4*>(p)[1] = make_float4(v[4], v[5], v[6], v[7]);
  }
  static __device__ __forceinline__ void store4(float* p, const float* v) {
    reinterpret_cast<float4*>(p)[0] = make_float4(v[0], v[1], v[2], v[3]);
  }
  static __device__ __forceinline__ void store16(float* p, const float* v) {  // 16 channels = two full sectors
    st_v8f(p, make_float4(v[0], v[1], v[2], v[3]), make_float4(v[4], v[5], v[6], v[7]));
    st_v8f(p + 8, make_float4(v[8], v[9], v[10], v[11]), make_float4(v[12], v[13], v[14], v[15]));
  }
};
template <> struct Act<bf16> {
  static __device__ __forceinline__ void load4(const bf16* p, float* v) {
    uint2 a = *reinterpret_cast<const uint2*>(p);
    __nv_bfloat162 lo = *reinterpret_cast<__nv_bfloat162*>(&a.x);
    __nv_bfloat162 hi = *reinterpret_cast<__nv_bfloat162*>(&a.y);
    float2 f0 = __bfloat1622float2(lo), f1 = __bfloat1622float2(hi);
    v[0] = f0.x; v[1] = f0.y; v[2] = f1.x; v[3] = f1.y;
  }
  static __device__ __forceinline__ uint32_t pack(float a, float b) {
    __nv_bfloat162 t = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&t);
  }
  static __device__ __forceinline__ void store8(bf16* p, const float* v) {
    uint4 u;
    u.x = pack(v[0], v[1]); u.y = pack(v[2], v[3]); u.z = pack(v[4], v[5]); u.w = pack(v[6], v[7]);
    *reinterpret_cast<uint4*>(p) = u;
  }
  static __device__ __forceinline__ void store4(bf16* p, const float* v) {
    uint2 u;
    u.x = pack(v[0], v[1]); u.y = pack(v[2], v[3]);
    *reinterpret_cast<uint2*>(p) = u;
  }
  static __device__ __forceinline__ void store16(bf16* p, const float* v) {  // 16 channels = one full 32-byte sector
    uint4 a, b;
    a.x = pack(v[0], v[1]); a.y = pack(v[2], v[3]); a.z = pack(v[4], v[5]); a.w = pack(v[6], v[7]);
    b.x = pack(v[8], v[9]); b.y = pack(v[10], v[11]); b.z = pack(v[12], v[13]); b.w = pack(v[14], v[15]);
    st_v8(p, a, b);
  }
};

struct Bil {
  int o00, o01, o10, o11;  // pixel offsets (y*W + x)
  float w00, w01, w10, w11;
};
// grid_sample(bilinear, border, align_corners=True) at pixel coordinate (fx, fy)
__device__ __forceinline__ Bil bil_setup(float fx, float fy, int W, int H) {
  fx = fminf(fmaxf(fx, 0.f), (float)(W - 1));
  fy = fminf(fmaxf(fy, 0.f), (float)(H - 1));
  float flx = floorf(fx), fly = floorf(fy);
  int x0 = (int)flx, y0 = (int)fly;
  int x1 = min(x0 + 1, W - 1), y1 = min(y0 + 1, H - 1);
  float wx = fx - flx, wy = fy - fly;
  Bil b;
  b.o00 = y0 * W + x0; b.o01 = y0 * W + x1; b.o10 = y1 * W + x0; b.o11 = y1 * W + x1;
  b.w00 = (1.f - wx) * (1.f - wy); b.w01 = wx * (1.f - wy); b.w10 = (1.f - wx) * wy; b.w11 = wx * wy;
  return b;
}
// byte k of a packed RGBX word as float, exact, on the full-rate pipes (PRMT + FADD instead of the quarter-rate I2F):
// 0x4B0000bb is the float 8388608 + bb
// PRMT takes ONE immediate: with the literal 0x4B000000 in it the selector travels through a register and every byte costs
// an extra MOV (a tenth of the blend kernel's instructions).  Read from constant memory the word lives in a register for the
// whole kernel and the selector becomes the immediate.
__constant__ uint32_t c_u8f_magic = 0x4B000000u;
template <int K>
__device__ __forceinline__ float u8f(uint32_t v) {
  return __uint_as_float(__byte_perm(v, c_u8f_magic, 0x7440 | K)) - 8388608.0f;
}
__device__ __forceinline__ void sample_rgbx(const uint8_t* img, const Bil& b, float* rgb) {
  const uint32_t* p = reinterpret_cast<const uint32_t*>(img);
  const uint32_t a = __ldg(p + b.o00), c = __ldg(p + b.o01), d = __ldg(p + b.o10), e = __ldg(p + b.o11);
  const float k = 1.0f / 255.0f;
  rgb[0] = (b.w00 * u8f<0>(a) + b.w01 * u8f<0>(c) + b.w10 * u8f<0>(d) + b.w11 * u8f<0>(e)) * k;
  rgb[1] = (b.w00 * u8f<1>(a) + b.w01 * u8f<1>(c) + b.w10 * u8f<1>(d) + b.w11 * u8f<1>(e)) * k;
  rgb[2] = (b.w00 * u8f<2>(a) + b.w01 * u8f<2>(c) + b.w10 * u8f<2>(d) + b.w11 * u8f<2>(e)) * k;
}
template <typename T>
__device__ __forceinline__ void sample_feat(const T* f, const Bil& b, float* o) {
  float a[4], c[4], d[4], e[4];
  Act<T>::load4(f + 4 * (size_t)b.o00, a);
  Act<T>::load4(f + 4 * (size_t)b.o01, c);
  Act<T>::load4(f + 4 * (size_t)b.o10, d);
  Act<T>::load4(f + 4 * (size_t)b.o11, e);
#pragma unroll
  for (int i = 0; i < 4; i++) o[i] = b.w00 * a[i] + b.w01 * c[i] + b.w10 * d[i] + b.w11 * e[i];
}

// The same three helpers on 32-bit PIXEL indices into the whole image set (base = image index * Hp * Wp folded into the
// offsets): one IMAD.WIDE per tap instead of 64-bit pointer arithmetic.  NI * Hp * Wp < 2^31 is checked by the host.
__device__ __forceinline__ Bil bil_setup(float fx, float fy, int W, int H, int base) {
  fx = fminf(fmaxf(fx, 0.f), (float)(W - 1));
  fy = fminf(fmaxf(fy, 0.f), (float)(H - 1));
  const int x0 = (int)fx, y0 = (int)fy;  // truncation == floor: both are >= 0
  const float wx = fx - (float)x0, wy = fy - (float)y0;
  const int dx = x0 < W - 1 ? 1 : 0, dy = y0 < H - 1 ? W : 0;
  Bil b;
  b.o00 = base + y0 * W + x0; b.o01 = b.o00 + dx; b.o10 = b.o00 + dy; b.o11 = b.o10 + dx;
  const float ux = 1.f - wx, uy = 1.f - wy;
  b.w00 = ux * uy; b.w01 = wx * uy; b.w10 = ux * wy; b.w11 = wx * wy;
  return b;
}
__device__ __forceinline__ void sample_rgbx(const uint32_t* __restrict__ imgs, const Bil& b, float* rgb) {
  const uint32_t a = __ldg(imgs + b.o00), c = __ldg(imgs + b.o01), d = __ldg(imgs + b.o10), e = __ldg(imgs + b.o11);
  const float k = 1.0f / 255.0f;
  rgb[0] = (b.w00 * u8f<0>(a) + b.w01 * u8f<0>(c) + b.w10 * u8f<0>(d) + b.w11 * u8f<0>(e)) * k;
  rgb[1] = (b.w00 * u8f<1>(a) + b.w01 * u8f<1>(c) + b.w10 * u8f<1>(d) + b.w11 * u8f<1>(e)) * k;
  rgb[2] = (b.w00 * u8f<2>(a) + b.w01 * u8f<2>(c) + b.w10 * u8f<2>(d) + b.w11 * u8f<2>(e)) * k;
}
__device__ __forceinline__ void feat4(const float* __restrict__ f, int pix, float* v) {
  const float4 a = __ldg(reinterpret_cast<const float4*>(f) + pix);
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
}
__device__ __forceinline__ void feat4(const bf16* __restrict__ f, int pix, float* v) {
  const uint2 a = __ldg(reinterpret_cast<const uint2*>(f) + pix);
  v[0] = __uint_as_float(a.x << 16); v[1] = __uint_as_float(a.x & 0xffff0000u);
  v[2] = __uint_as_float(a.y << 16); v[3] = __uint_as_float(a.y & 0xffff0000u);
}
template <typename T>
__device__ __forceinline__ void sample_feat_px(const T* __restrict__ feats, const Bil& b, float* o) {
  float a[4], c[4], d[4], e[4];
  feat4(feats, b.o00, a); feat4(feats, b.o01, c); feat4(feats, b.o10, d); feat4(feats, b.o11, e);
#pragma unroll
  for (int i = 0; i < 4; i++) o[i] = b.w00 * a[i] + b.w01 * c[i] + b.w10 * d[i] + b.w11 * e[i];
}

// ------------------------------------------------------------------------------------------
// to_tensor + pad_to_multiple(reflect) (models/base.py:175-186, :203-225) kept in uint8: RGB -> RGBX
// edge: 0 = torch 'reflect' / cv2 BORDER_REFLECT_101 (models/base.py:203-225), 1 = cv2.BORDER_REFLECT, the border pixel
// repeated (scripts/run_experiments.py:377-380); bgr: the caller's frames are BGR (run_experiments.py:381-382 flips them
// for the network): the padded copy is RGB either way
__global__ void pad_rgbx_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, int B, int H, int W,
                                int Hp, int Wp, int edge, int bgr) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  size_t total = (size_t)B * Hp * Wp;
  if (i >= total) return;
  int x = (int)(i % Wp);
  int y = (int)((i / Wp) % Hp);
  int b = (int)(i / ((size_t)Wp * Hp));
  int sx = x < W ? x : 2 * (W - 1) + edge - x;
  int sy = y < H ? y : 2 * (H - 1) + edge - y;
  const uint8_t* p = src + (((size_t)b * H + sy) * W + sx) * 3;
  reinterpret_cast<uchar4*>(dst)[i] = bgr ? make_uchar4(p[2], p[1], p[0], 0) : make_uchar4(p[0], p[1], p[2], 0);
}
// the same for W % 4 == 0 and a 4-byte aligned source: a thread turns 12 source bytes (three aligned words) into four
// RGBX pixels (one 16-byte store); only the reflected border falls back to byte loads
__global__ void pad_rgbx4_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, int B, int H, int W, int Hp,
                                 int Wp, int edge, int bgr) {
  const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  const int Wq = Wp >> 2;
  if (i >= (size_t)B * Hp * Wq) return;
  const int x = (int)(i % Wq) * 4, y = (int)((i / Wq) % Hp), b = (int)(i / ((size_t)Wq * Hp));
  const int sy = y < H ? y : 2 * (H - 1) + edge - y;
  const uint8_t* row = src + ((size_t)b * H + sy) * W * 3;
  uint4 o;
  if (x + 3 < W) {
    const uint32_t* p = reinterpret_cast<const uint32_t*>(row + (size_t)x * 3);
    const uint32_t w0 = __ldg(p), w1 = __ldg(p + 1), w2 = __ldg(p + 2);
    o.x = w0 & 0x00ffffffu;
    o.y = (w0 >> 24) | ((w1 & 0xffffu) << 8);
    o.z = (w1 >> 16) | ((w2 & 0xffu) << 16);
    o.w = w2 >> 8;
  } else {
    uint32_t v[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const int xx = x + k, sx = xx < W ? xx : 2 * (W - 1) + edge - xx;
      const uint8_t* p = row + (size_t)sx * 3;
      v[k] = (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16);
    }
    o = make_uint4(v[0], v[1], v[2], v[3]);
  }
  if (bgr) { o.x = __byte_perm(o.x, 0, 0x4012); o.y = __byte_perm(o.y, 0, 0x4012); o.z = __byte_perm(o.z, 0, 0x4012); o.w = __byte_perm(o.w, 0, 0x4012); }
  reinterpret_cast<uint4*>(dst)[i] = o;
}
int k_pad_rgbx(vfi_handle* h, const uint8_t* src, uint8_t* dst, int B, int H, int W, int Hp, int Wp, cudaStream_t s, int edge, int bgr) {
  ProfScope _ps(h, PC_PAD, s);
  size_t total = (size_t)B * Hp * Wp;
  if ((W & 3) == 0 && (Wp & 3) == 0 && (reinterpret_cast<uintptr_t>(src) & 3) == 0)
    pad_rgbx4_kernel<<<cdiv(total / 4, 256), 256, 0, s>>>(src, dst, B, H, W, Hp, Wp, edge, bgr);
  else
    pad_rgbx_kernel<<<cdiv(total, 256), 256, 0, s>>>(src, dst, B, H, W, Hp, Wp, edge, bgr);
  VFI_LAUNCHED(h);
  return 0;
}

// ------------------------------------------------------------------------------------------
// Head.cnn0: Conv2d(3,16,3,2,1) + LeakyReLU straight from the uint8 image (K = 27: CUDA cores).  The 432 weights
// (pre-divided by 255) and the bias travel as a kernel parameter, so every FFMA takes its weight as a constant-bank
// operand: no weight loads at all, the kernel is bound by its 432 FMAs per output pixel.
struct Head0W {
  float w[27 * 16];
  float b[16];
};
template <typename T>
__global__ void __launch_bounds__(128) head_cnn0_kernel(const uint8_t* __restrict__ imgs, const __grid_constant__ Head0W Wt,
                                                        T* __restrict__ out, int NI, int Hp, int Wp, float slope) {
  int Ho = Hp / 2, Wo = Wp / 2;
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  size_t total = (size_t)NI * Ho * Wo;
  if (i >= total) return;
  int x = (int)(i % Wo), y = (int)((i / Wo) % Ho), n = (int)(i / ((size_t)Wo * Ho));
  const uint32_t* img = reinterpret_cast<const uint32_t*>(imgs) + (size_t)n * Hp * Wp;
  float acc[16];
#pragma unroll
  for (int c = 0; c < 16; c++) acc[c] = Wt.b[c];
#pragma unroll
  for (int r = 0; r < 3; r++) {
    const int iy = 2 * y + r - 1;
#pragma unroll
    for (int q = 0; q < 3; q++) {
      const int ix = 2 * x + q - 1;
      // zero padding: only the top row / left column can fall outside (Hp, Wp are even)
      const bool in = (r > 0 || iy >= 0) && (q > 0 || ix >= 0);
      const uint32_t p = in ? __ldg(img + (size_t)iy * Wp + ix) : 0u;
      const float v[3] = {u8f<0>(p), u8f<1>(p), u8f<2>(p)};
#pragma unroll
      for (int ci = 0; ci < 3; ci++)
#pragma unroll
        for (int c = 0; c < 16; c++) acc[c] = fmaf(v[ci], Wt.w[((r * 3 + q) * 3 + ci) * 16 + c], acc[c]);
    }
  }
#pragma unroll
  for (int c = 0; c < 16; c++) acc[c] = acc[c] > 0.f ? acc[c] : acc[c] * slope;
  T* o = out + i * 16;
  Act<T>::store16(o, acc);
}
int k_head_cnn0(vfi_handle* h, const Layer& L, const uint8_t* imgs, void* out, int NI, int Hp, int Wp, bool bf,
                cudaStream_t s) {
  ProfScope _ps(h, PC_HEAD0, s);
  (void)L;
  size_t total = (size_t)NI * (Hp / 2) * (Wp / 2);
  Head0W Wt;
  memcpy(Wt.w, h->head0_w, sizeof(Wt.w));
  memcpy(Wt.b, h->head0_b, sizeof(Wt.b));
  if (bf) head_cnn0_kernel<bf16><<<cdiv(total, 128), 128, 0, s>>>(imgs, Wt, (bf16*)out, NI, Hp, Wp, h->slope);
  else head_cnn0_kernel<float><<<cdiv(total, 128), 128, 0, s>>>(imgs, Wt, (float*)out, NI, Hp, Wp, h->slope);
  VFI_LAUNCHED(h);
  return 0;
}

// ------------------------------------------------------------------------------------------
// F.interpolate(scale_factor=s, bilinear, align_corners=False) of a coarse tensor, evaluated at ONE full-resolution
// pixel: area_pixel_compute_source_index(scale = 1/s) = max(0, (dst + 0.5)/s - 0.5)   
struct Lerp {
  int o00, o01, o10, o11;  // texel offsets (y*Wl + x) inside one image of the coarse tensor
  float hx, lx, hy, ly;
};
__device__ __forceinline__ Lerp lerp_setup(int x, int y, int s, int Hl, int Wl) {
  const float rs = 1.0f / (float)s;
  float sy = fmaxf(rs * ((float)y + 0.5f) - 0.5f, 0.f);
  float sx = fmaxf(rs * ((float)x + 0.5f) - 0.5f, 0.f);
  int y0 = (int)sy, x0 = (int)sx;
  int y1 = y0 + (y0 < Hl - 1 ? 1 : 0), x1 = x0 + (x0 < Wl - 1 ? 1 : 0);
  Lerp L;
  L.ly = sy - (float)y0; L.lx = sx - (float)x0;
  L.hy = 1.f - L.ly; L.hx = 1.f - L.lx;
  L.o00 = y0 * Wl + x0; L.o01 = y0 * Wl + x1; L.o10 = y1 * Wl + x0; L.o11 = y1 * Wl + x1;
  return L;
}
// channels 4k..4k+3 of a coarse tensor with c4 float4 per texel
__device__ __forceinline__ float4 lerp4(const float4* img, const Lerp& L, int k, int c4) {
  float4 a = __ldg(img + c4 * L.o00 + k), b = __ldg(img + c4 * L.o01 + k), c = __ldg(img + c4 * L.o10 + k),
         d = __ldg(img + c4 * L.o11 + k);
  float4 r;
  r.x = L.hy * (L.hx * a.x + L.lx * b.x) + L.ly * (L.hx * c.x + L.lx * d.x);
  r.y = L.hy * (L.hx * a.y + L.lx * b.y) + L.ly * (L.hx * c.y + L.lx * d.y);
  r.z = L.hy * (L.hx * a.z + L.lx * b.z) + L.ly * (L.hx * c.z + L.lx * d.z);
  r.w = L.hy * (L.hx * a.w + L.lx * b.w) + L.ly * (L.hx * c.w + L.lx * d.w);
  return r;
}
// The same source index / weights for every column and row of the padded frame, once per call: the tiled kernels read
// them instead of redoing the float arithmetic per thread and level.
struct TapScales { int s[5]; };
__global__ void build_taps_kernel(Tap* __restrict__ out, const TapScales sc, int n, int Hp, int Wp) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x, k = blockIdx.y;
  if (i >= Wp + Hp || k >= n) return;
  const int s = sc.s[k];
  const int d = i < Wp ? i : i - Wp, L = (i < Wp ? Wp : Hp) / s;
  const float rs = 1.0f / (float)s;
  const float src = fmaxf(rs * ((float)d + 0.5f) - 0.5f, 0.f);
  Tap t;
  t.i0 = (int)src; t.i1 = t.i0 + (t.i0 < L - 1 ? 1 : 0);
  t.l = src - (float)t.i0; t.h = 1.f - t.l;
  out[(size_t)k * (Wp + Hp) + i] = t;
}
int k_build_taps(vfi_handle* h, Tap* out, const int* scales, int n, int Hp, int Wp, cudaStream_t s) {
  ProfScope _ps(h, PC_MISC, s);
  TapScales sc;
  for (int k = 0; k < 5; k++) sc.s[k] = k < n ? scales[k] : 1;
  build_taps_kernel<<<dim3(cdiv(Wp + Hp, 256), n), 256, 0, s>>>(out, sc, n, Hp, Wp);
  VFI_LAUNCHED(h);
  return 0;
}

// mask + feat 0..7 of texel `idx` of a level stored as LowPlanes (bf16 -> fp32 is a shift)
__device__ __forceinline__ void mf9_load(const bf16* __restrict__ mf, const bf16* __restrict__ f7, int idx, float* v) {
  const uint4 a = __ldg(reinterpret_cast<const uint4*>(mf) + idx);
  const uint32_t h = __ldg(reinterpret_cast<const unsigned short*>(f7) + idx);
  v[0] = __uint_as_float(a.x << 16); v[1] = __uint_as_float(a.x & 0xffff0000u);
  v[2] = __uint_as_float(a.y << 16); v[3] = __uint_as_float(a.y & 0xffff0000u);
  v[4] = __uint_as_float(a.z << 16); v[5] = __uint_as_float(a.z & 0xffff0000u);
  v[6] = __uint_as_float(a.w << 16); v[7] = __uint_as_float(a.w & 0xffff0000u);
  v[8] = __uint_as_float(h << 16);
}
// the MF = 1 part of state_at for a LowPlanes level: m[0] = (mask, f0, f1, f2), m[1] = (f3..f6), m[2].x = f7
__device__ __forceinline__ void mf9_lerp(const bf16* __restrict__ mf, const bf16* __restrict__ f7, int tb, const Lerp& L,
                                         float4* m) {
  float a[9], b[9], c[9], d[9], r[9];
  mf9_load(mf, f7, tb + L.o00, a); mf9_load(mf, f7, tb + L.o01, b);
  mf9_load(mf, f7, tb + L.o10, c); mf9_load(mf, f7, tb + L.o11, d);
#pragma unroll
  for (int i = 0; i < 9; i++) r[i] = L.hy * (L.hx * a[i] + L.lx * b[i]) + L.ly * (L.hx * c[i] + L.lx * d[i]);
  m[0] = make_float4(r[0], r[1], r[2], r[3]); m[1] = make_float4(r[4], r[5], r[6], r[7]);
  m[2] = make_float4(r[8], 0.f, 0.f, 0.f);
}
__global__ void unpack_planes_kernel(const float4* __restrict__ flow, const bf16* __restrict__ mf, const bf16* __restrict__ f7,
                                     float4* __restrict__ out, size_t texels) {
  const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i >= texels) return;
  float v[9];
  mf9_load(mf, f7, (int)i, v);
  out[4 * i] = flow[i];
  out[4 * i + 1] = make_float4(v[0], v[1], v[2], v[3]);
  out[4 * i + 2] = make_float4(v[4], v[5], v[6], v[7]);
  out[4 * i + 3] = make_float4(v[8], 0.f, 0.f, 0.f);
}
// test hook: a LowPlanes level as the fp32 [texels, 16] layout of the other precision
int k_unpack_planes(vfi_handle* h, const LowPlanes& pl, float* out16, size_t texels, cudaStream_t s) {
  ProfScope _ps(h, PC_MISC, s);
  unpack_planes_kernel<<<cdiv(texels, 256), 256, 0, s>>>(reinterpret_cast<const float4*>(pl.flow), pl.mf, pl.f7,
                                                         reinterpret_cast<float4*>(out16), texels);
  VFI_LAUNCHED(h);
  return 0;
}

// flow after the blocks in `lows` at full-resolution pixel (x,y) of item n: flow = flow + up(low_k)[:4] * s_k, in
// block order.  MF = 1: also (mask, feat) = up(low_last)[4:13] in m[0..2];  MF = 2: mask only, in m[0].x.
// A low at scale 1 (the last block at scale=1.0) is read directly: interpolate(x1) is the identity.
template <int MF, bool PL = false>
__device__ __forceinline__ float4 state_at(const LowSet& lows, int n, int x, int y, int Hp, int Wp, float4* m) {
  float4 fl = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
  for (int k = 0; k < 5; k++) {
    if (k < lows.n) {
      const int sk = lows.s[k], c4 = lows.c4[k], Hl = Hp / sk, Wl = Wp / sk;
      const float4* img = reinterpret_cast<const float4*>(lows.p[k]) + (size_t)n * Hl * Wl * c4;
      const bool last = k == lows.n - 1;
      if (sk == 1) {
        const float4* px = img + (size_t)(y * Wl + x) * c4;
        const float4 r = __ldg(px);
        fl.x += r.x; fl.y += r.y; fl.z += r.z; fl.w += r.w;
        if (MF == 1 && last) { m[0] = __ldg(px + 1); m[1] = __ldg(px + 2); m[2] = __ldg(px + 3); }
        if (MF == 2 && last) m[0] = __ldg(px + 1);
      } else {
        const Lerp L = lerp_setup(x, y, sk, Hl, Wl);
        const float4 r = lerp4(img, L, 0, c4);
        const float fs = (float)sk;
        fl.x += r.x * fs; fl.y += r.y * fs; fl.z += r.z * fs; fl.w += r.w * fs;
        if (MF == 1 && last) {
          if (PL) mf9_lerp(lows.mf[k], lows.f7[k], n * Hl * Wl, L, m);
          else { m[0] = lerp4(img, L, 1, c4); m[1] = lerp4(img, L, 2, c4); m[2] = lerp4(img, L, 3, c4); }
        }
        if (MF == 2 && last) m[0] = lerp4(img, L, 1, c4);
      }
    }
  }
  return fl;
}

// ------------------------------------------------------------------------------------------
// Block-input gather.  Produces IFBlock's conv0 input at 1/s resolution in ONE pass:
//   block 0 : cat(img0, img1, f0, f1, t)                               -> 15 ch (+1 zero)
//   block i : cat(warp(img0,F01), warp(img1,F10), warp(f0,F01), warp(f1,F10), t, mask, feat, flow/s)
//                                                                       -> 28 ch (+4 zero)
// followed by F.interpolate(1/s, bilinear, align_corners=False), which for s in {2,4,8,16,32}
// is exactly the mean of the 2x2 full-resolution pixels at offsets s/2-1, s/2,
// so only those pixels are ever warped, and flow / mask / feat are evaluated only there, straight from the
// coarse lastconv outputs of the previous blocks (state_at).  Neither the 28-channel full-resolution concat
// nor the full-resolution flow of the reference ever exists.  Output NHWC, plain or split into the 4 stride-2
// phase planes the conv0a TMA reads.
template <typename T, bool FIRST>
__global__ void gather_kernel(const uint8_t* __restrict__ imgs, const T* __restrict__ feats,
                              const __grid_constant__ LowSet lows, const float* __restrict__ tsteps,
                              T* __restrict__ out, int B, int I1, int Tn, int Hp, int Wp, int s, int phase) {
  constexpr int C = FIRST ? 16 : 32;
  const int Hs = Hp / s, Ws = Wp / s;
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  size_t total = (size_t)B * Tn * Hs * Ws;
  if (i >= total) return;
  int X = (int)(i % Ws), Y = (int)((i / Ws) % Hs), n = (int)(i / ((size_t)Ws * Hs));
  int b = n / Tn;
  float t = tsteps[n % Tn];
  const size_t plane = (size_t)Hp * Wp;
  const uint8_t* img0 = imgs + ((size_t)b) * plane * 4;
  const uint8_t* img1 = imgs + ((size_t)(I1 + b)) * plane * 4;
  const T* f0 = feats + ((size_t)b) * plane * 4;
  const T* f1 = feats + ((size_t)(I1 + b)) * plane * 4;
  const int ns = s == 1 ? 1 : 2;
  const int by = s == 1 ? Y : s * Y + s / 2 - 1;
  const int bx = s == 1 ? X : s * X + s / 2 - 1;
  float acc[C];
#pragma unroll
  for (int c = 0; c < C; c++) acc[c] = 0.f;
  for (int dy = 0; dy < ns; dy++)
    for (int dx = 0; dx < ns; dx++) {
      int y = by + dy, x = bx + dx;
      float v[4];
      if constexpr (FIRST) {
        uchar4 p0 = __ldg(reinterpret_cast<const uchar4*>(img0) + (size_t)y * Wp + x);
        uchar4 p1 = __ldg(reinterpret_cast<const uchar4*>(img1) + (size_t)y * Wp + x);
        acc[0] += p0.x / 255.0f; acc[1] += p0.y / 255.0f; acc[2] += p0.z / 255.0f;
        acc[3] += p1.x / 255.0f; acc[4] += p1.y / 255.0f; acc[5] += p1.z / 255.0f;
        Act<T>::load4(f0 + 4 * ((size_t)y * Wp + x), v);
        acc[6] += v[0]; acc[7] += v[1]; acc[8] += v[2]; acc[9] += v[3];
        Act<T>::load4(f1 + 4 * ((size_t)y * Wp + x), v);
        acc[10] += v[0]; acc[11] += v[1]; acc[12] += v[2]; acc[13] += v[3];
      } else {
        float4 m[3];
        const float4 fl = state_at<1, std::is_same<T, bf16>::value>(lows, n, x, y, Hp, Wp, m);
        const float4 s1 = m[0], s2 = m[1], s3 = m[2];
        Bil b0 = bil_setup((float)x + fl.x, (float)y + fl.y, Wp, Hp);
        Bil b1 = bil_setup((float)x + fl.z, (float)y + fl.w, Wp, Hp);
        float rgb[3];
        sample_rgbx(img0, b0, rgb);
        acc[0] += rgb[0]; acc[1] += rgb[1]; acc[2] += rgb[2];
        sample_rgbx(img1, b1, rgb);
        acc[3] += rgb[0]; acc[4] += rgb[1]; acc[5] += rgb[2];
        sample_feat<T>(f0, b0, v);
        acc[6] += v[0]; acc[7] += v[1]; acc[8] += v[2]; acc[9] += v[3];
        sample_feat<T>(f1, b1, v);
        acc[10] += v[0]; acc[11] += v[1]; acc[12] += v[2]; acc[13] += v[3];
        acc[15] += s1.x;                                                 // mask (pre-sigmoid)
        acc[16] += s1.y; acc[17] += s1.z; acc[18] += s1.w;              // feat 0..2
        acc[19] += s2.x; acc[20] += s2.y; acc[21] += s2.z; acc[22] += s2.w;
        acc[23] += s3.x;
        acc[24] += fl.x; acc[25] += fl.y; acc[26] += fl.z; acc[27] += fl.w;
      }
    }
  if (ns == 2) {
#pragma unroll
    for (int c = 0; c < C; c++) acc[c] *= 0.25f;
  }
  acc[14] = t;
  if constexpr (!FIRST) {
    const float inv = 1.0f / (float)s;
    acc[24] *= inv; acc[25] *= inv; acc[26] *= inv; acc[27] *= inv;
  }
  size_t o;
  if (phase) {
    int H2 = Hs / 2, W2 = Ws / 2;
    o = ((((size_t)n * 4 + (Y & 1) * 2 + (X & 1)) * H2 + (Y >> 1)) * W2 + (X >> 1)) * C;
  } else {
    o = i * C;
  }
#pragma unroll
  for (int c = 0; c < C; c += 16) Act<T>::store16(out + o + c, acc + c);
}

// ------------------------------------------------------------------------------------------
// Tiled evaluation of the same state for kernels that visit EVERY full-resolution pixel (the gathers of the two
// finest blocks and the final blend).  A CTA owns a 32x32 pixel tile.  Stage 1 writes, for every coarse level, the
// x-interpolated rows that cover the tile into shared memory: T[row][x] = hx*a + lx*b -- exactly the inner terms of
// bilinear upsampling, out = hy*(hx*a + lx*b) + ly*(hx*c + lx*d).  Stage 2 finishes a pixel with two conflict-free
// LDS.128 per level (hy*T[y0][x] + ly*T[y1][x]): the arithmetic, and so the result, is bit-identical to state_at,
// with ~1.2 global texel loads per pixel instead of 16 per level.
#define TL_W 32
#define TL_H 32
#define TL_THREADS 256
#define TL_TAB (5 * TL_H)  // float4 slots of the row table at the start of the tile's shared memory
// per (level, tile row): where the two staged coarse rows of that pixel row start (float4 offsets) + the y weights
struct __align__(16) RowTap {
  int o0, o1;
  float hy, ly;
};
__host__ __device__ __forceinline__ int tl_rows(int s) { return (TL_H / s > 1 ? TL_H / s : 1) + 2; }
__host__ __device__ __forceinline__ int tl_nv(int mf) { return mf == 1 ? 4 : mf == 2 ? 2 : 1; }
static size_t tl_smem_bytes(const LowSet& lows, int mf) {
  size_t n = TL_TAB;
  for (int k = 0; k < lows.n; k++) {
    if (lows.s[k] == 1) continue;
    n += (size_t)tl_rows(lows.s[k]) * ((k == lows.n - 1) ? tl_nv(mf) : 1) * TL_W;
  }
  return n * sizeof(float4);
}
// Stage 1.  Thread t owns tile column t%32 (its x taps are computed once per level) and rows t/32, t/32+8, ...;
// warp k also fills level k's row table.
template <int MF, int NT = TL_THREADS, bool PL = false>
__device__ __forceinline__ void tile_stage(const LowSet& L, int n, int tx0, int ty0, int Hp, int Wp, float4* sm) {
  RowTap* ytab = reinterpret_cast<RowTap*>(sm);
  const int xl = threadIdx.x & (TL_W - 1), row0 = threadIdx.x >> 5;
  int off = TL_TAB;
#pragma unroll
  for (int k = 0; k < 5; k++) {
    if (k < L.n && L.s[k] > 1) {
      const int sk = L.s[k], c4 = L.c4[k], Hl = L.Hl[k], Wl = L.Wl[k];
      const int nv = (k == L.n - 1) ? tl_nv(MF) : 1;
      const int rows = tl_rows(sk);
      const float rs = L.rs[k];
      const int r_lo = (int)fmaxf(rs * ((float)ty0 + 0.5f) - 0.5f, 0.f);
      if (row0 == k) {
        const float sy = fmaxf(rs * ((float)min(ty0 + xl, Hp - 1) + 0.5f) - 0.5f, 0.f);
        const int y0 = (int)sy, y1 = y0 + (y0 < Hl - 1 ? 1 : 0);
        RowTap t;
        t.ly = sy - (float)y0; t.hy = 1.f - t.ly;
        t.o0 = off + (y0 - r_lo) * nv * TL_W; t.o1 = off + (y1 - r_lo) * nv * TL_W;
        ytab[k * TL_H + xl] = t;
      }
      const float sx = fmaxf(rs * ((float)min(tx0 + xl, Wp - 1) + 0.5f) - 0.5f, 0.f);
      const int x0 = (int)sx, x1 = x0 + (x0 < Wl - 1 ? 1 : 0);
      const float lx = sx - (float)x0, hx = 1.f - lx;
      const float4* img = reinterpret_cast<const float4*>(L.p[k]) + (size_t)n * Hl * Wl * c4;
      for (int row = row0; row < rows; row += NT / TL_W) {
        const int ry = min(r_lo + row, Hl - 1);
        const float4* pa = img + (size_t)(ry * Wl + x0) * c4;
        const float4* pb = img + (size_t)(ry * Wl + x1) * c4;
        float4* dst = sm + off + row * nv * TL_W + xl;
        if (PL && MF == 1 && k == L.n - 1) {  // LowPlanes level: flow from its fp32 plane, mask / feat from the bf16 ones
          const float4 a = __ldg(pa), b = __ldg(pb);
          dst[0] = make_float4(hx * a.x + lx * b.x, hx * a.y + lx * b.y, hx * a.z + lx * b.z, hx * a.w + lx * b.w);
          const int tb = (n * Hl + ry) * Wl;
          float u[9], w[9];
          mf9_load(L.mf[k], L.f7[k], tb + x0, u);
          mf9_load(L.mf[k], L.f7[k], tb + x1, w);
#pragma unroll
          for (int i = 0; i < 9; i++) u[i] = hx * u[i] + lx * w[i];
          dst[TL_W] = make_float4(u[0], u[1], u[2], u[3]);
          dst[2 * TL_W] = make_float4(u[4], u[5], u[6], u[7]);
          dst[3 * TL_W] = make_float4(u[8], 0.f, 0.f, 0.f);
          continue;
        }
#pragma unroll
        for (int v = 0; v < 4; v++) {
          if (v < nv) {
            const float4 a = __ldg(pa + v), b = __ldg(pb + v);
            dst[v * TL_W] =
                make_float4(hx * a.x + lx * b.x, hx * a.y + lx * b.y, hx * a.z + lx * b.z, hx * a.w + lx * b.w);
          }
        }
      }
      off += rows * nv * TL_W;
    }
  }
}
// Stage 2: state_at for pixel (x,y) = tile-local (xl,yl)
template <int MF>
__device__ __forceinline__ float4 tile_eval(const LowSet& L, const float4* sm, int n, int x, int y, int xl, int yl,
                                            float4* m) {
  const RowTap* ytab = reinterpret_cast<const RowTap*>(sm);
  float4 fl = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
  for (int k = 0; k < 5; k++) {
    if (k < L.n) {
      const bool last = k == L.n - 1;
      if (L.s[k] == 1) {
        const int c4 = L.c4[k];
        const float4* px = reinterpret_cast<const float4*>(L.p[k]) + ((size_t)n * L.Hl[k] + y) * L.Wl[k] * c4 + x * c4;
        const float4 r = __ldg(px);
        fl.x += r.x; fl.y += r.y; fl.z += r.z; fl.w += r.w;
        if (MF == 1 && last) { m[0] = __ldg(px + 1); m[1] = __ldg(px + 2); m[2] = __ldg(px + 3); }
        if (MF == 2 && last) m[0] = __ldg(px + 1);
      } else {
        const RowTap t = ytab[k * TL_H + yl];
        const float4* r0 = sm + t.o0 + xl;
        const float4* r1 = sm + t.o1 + xl;
        const float4 a = r0[0], c = r1[0];
        const float fs = L.fs[k];
        fl.x += (t.hy * a.x + t.ly * c.x) * fs; fl.y += (t.hy * a.y + t.ly * c.y) * fs;
        fl.z += (t.hy * a.z + t.ly * c.z) * fs; fl.w += (t.hy * a.w + t.ly * c.w) * fs;
        if (MF != 0 && last) {
#pragma unroll
          for (int v = 1; v < tl_nv(MF); v++) {
            const float4 p = r0[v * TL_W], q = r1[v * TL_W];
            m[v - 1] = make_float4(t.hy * p.x + t.ly * q.x, t.hy * p.y + t.ly * q.y, t.hy * p.z + t.ly * q.z,
                                   t.hy * p.w + t.ly * q.w);
          }
        }
      }
    }
  }
  return fl;
}

// one warped tap of a block input (channels as in gather_kernel), accumulated into acc[28]
template <typename T>
__device__ __forceinline__ void gather_tap(const uint8_t* img0, const uint8_t* img1, const T* f0, const T* f1,
                                           const float4 fl, const float4* m, int x, int y, int Hp, int Wp,
                                           float* acc) {
  Bil b0 = bil_setup((float)x + fl.x, (float)y + fl.y, Wp, Hp);
  Bil b1 = bil_setup((float)x + fl.z, (float)y + fl.w, Wp, Hp);
  float rgb[3], v[4];
  sample_rgbx(img0, b0, rgb);
  acc[0] += rgb[0]; acc[1] += rgb[1]; acc[2] += rgb[2];
  sample_rgbx(img1, b1, rgb);
  acc[3] += rgb[0]; acc[4] += rgb[1]; acc[5] += rgb[2];
  sample_feat<T>(f0, b0, v);
  acc[6] += v[0]; acc[7] += v[1]; acc[8] += v[2]; acc[9] += v[3];
  sample_feat<T>(f1, b1, v);
  acc[10] += v[0]; acc[11] += v[1]; acc[12] += v[2]; acc[13] += v[3];
  acc[15] += m[0].x;
  acc[16] += m[0].y; acc[17] += m[0].z; acc[18] += m[0].w;
  acc[19] += m[1].x; acc[20] += m[1].y; acc[21] += m[1].z; acc[22] += m[1].w;
  acc[23] += m[2].x;
  acc[24] += fl.x; acc[25] += fl.y; acc[26] += fl.z; acc[27] += fl.w;
}

// gather_kernel for the blocks at s = 1 and s = 2, where the taps cover every full-resolution pixel.
// grid = (Wp/32, Hp/32, N)
template <typename T, int S>
__global__ void __launch_bounds__(TL_THREADS) gather_tile_kernel(const uint8_t* __restrict__ imgs,
                                                                 const T* __restrict__ feats,
                                                                 const __grid_constant__ LowSet lows,
                                                                 const float* __restrict__ tsteps,
                                                                 T* __restrict__ out, int B, int I1, int Tn, int Hp, int Wp,
                                                                 int phase) {
  extern __shared__ float4 tl_sm[];
  const int n = blockIdx.z, tx0 = blockIdx.x * TL_W, ty0 = blockIdx.y * TL_H;
  tile_stage<1, TL_THREADS, std::is_same<T, bf16>::value>(lows, n, tx0, ty0, Hp, Wp, tl_sm);
  __syncthreads();
  const int b = n / Tn;
  const float t = tsteps[n - b * Tn];
  const size_t plane = (size_t)Hp * Wp;
  const uint8_t* img0 = imgs + ((size_t)b) * plane * 4;
  const uint8_t* img1 = imgs + ((size_t)(I1 + b)) * plane * 4;
  const T* f0 = feats + ((size_t)b) * plane * 4;
  const T* f1 = feats + ((size_t)(I1 + b)) * plane * 4;
  const int Hs = Hp / S, Ws = Wp / S;
  constexpr int NOUT = S == 1 ? 4 : 1;  // outputs per thread
#pragma unroll 2
  for (int i = 0; i < NOUT; i++) {
    float acc[32];
#pragma unroll
    for (int c = 0; c < 32; c++) acc[c] = 0.f;
    int X, Y;
    if (S == 1) {
      const int xl = threadIdx.x & (TL_W - 1), yl = (threadIdx.x >> 5) + (TL_THREADS / TL_W) * i;
      X = tx0 + xl; Y = ty0 + yl;
      float4 m[3];
      const float4 fl = tile_eval<1>(lows, tl_sm, n, X, Y, xl, yl, m);
      gather_tap<T>(img0, img1, f0, f1, fl, m, X, Y, Hp, Wp, acc);
    } else {
      const int Xl = threadIdx.x & (TL_W / 2 - 1), Yl = threadIdx.x >> 4;
      X = (tx0 >> 1) + Xl; Y = (ty0 >> 1) + Yl;
#pragma unroll 2
      for (int d = 0; d < 4; d++) {
        // odd groups of four outputs visit their right-hand tap first: the 8 lanes of an LDS.128 phase then hit 8
        // distinct 16-byte bank groups
        const int xl = 2 * Xl + ((d & 1) ^ ((Xl >> 2) & 1)), yl = 2 * Yl + (d >> 1);
        float4 m[3];
        const float4 fl = tile_eval<1>(lows, tl_sm, n, tx0 + xl, ty0 + yl, xl, yl, m);
        gather_tap<T>(img0, img1, f0, f1, fl, m, tx0 + xl, ty0 + yl, Hp, Wp, acc);
      }
#pragma unroll
      for (int c = 0; c < 32; c++) acc[c] *= 0.25f;
    }
    acc[14] = t;
    const float inv = 1.0f / (float)S;
    acc[24] *= inv; acc[25] *= inv; acc[26] *= inv; acc[27] *= inv;
    size_t o;
    if (phase) {
      const int H2 = Hs / 2, W2 = Ws / 2;
      o = ((((size_t)n * 4 + (Y & 1) * 2 + (X & 1)) * H2 + (Y >> 1)) * W2 + (X >> 1)) * 32;
    } else {
      o = (((size_t)n * Hs + Y) * Ws + X) * 32;
    }
#pragma unroll
    for (int c = 0; c < 32; c += 16) Act<T>::store16(out + o + c, acc + c);
  }
}

static int launch_gather_s2(vfi_handle* h, const uint8_t* imgs, const bf16* feats, const LowSet& lows, const float* tsteps_dev,
                            bf16* out, int B, int I1, int T, int Hp, int Wp, bool phase, cudaStream_t s);
static int launch_gather_lo(vfi_handle* h, const uint8_t* imgs, const bf16* feats, const LowSet& lows, const float* tsteps_dev,
                            bf16* out, int B, int I1, int T, int Hp, int Wp, int s_int, bool phase, cudaStream_t s);
int k_gather(vfi_handle* h, const uint8_t* imgs, const void* feats, const LowSet& lows, const float* tsteps_dev,
             void* out, int B, int I1, int T, int Hp, int Wp, int s_int, bool bf, bool phase, cudaStream_t s) {
  ProfScope _ps(h, PC_GATHER, s);
  size_t total = (size_t)B * T * (Hp / s_int) * (Wp / s_int);
  int grid = cdiv(total, 128);
  const bool first = lows.n == 0;
  if (!first && bf && s_int == 2 && Hp % TL_H == 0 && Wp % TL_W == 0 && lows.mf[lows.n - 1] && lows.xt[0] &&
      lows.s[lows.n - 1] >= 4) {
    bool coarse = true;  // flat staging pre-supposes every level is coarser than the frame
    for (int k = 0; k < lows.n; k++) coarse = coarse && lows.s[k] > 1;
    if (coarse) return launch_gather_s2(h, imgs, (const bf16*)feats, lows, tsteps_dev, (bf16*)out, B, I1, T, Hp, Wp, phase, s);
  }
  if (!first && bf && s_int >= 4 && lows.xt[0] && lows.s[lows.n - 1] >= 2 * s_int) {
    bool planes = true;
    for (int k = 0; k < lows.n; k++) planes = planes && lows.mf[k] != nullptr;
    if (planes) return launch_gather_lo(h, imgs, (const bf16*)feats, lows, tsteps_dev, (bf16*)out, B, I1, T, Hp, Wp, s_int, phase, s);
  }
  if (!first && s_int <= 2 && Hp % TL_H == 0 && Wp % TL_W == 0) {
    const size_t smem = tl_smem_bytes(lows, 1);
    const dim3 tgrid(Wp / TL_W, Hp / TL_H, B * T);
#define GT_LAUNCH(TT, SS)                                                                                          \
  do {                                                                                                             \
    VFI_CUDA(h, cudaFuncSetAttribute(gather_tile_kernel<TT, SS>, cudaFuncAttributeMaxDynamicSharedMemorySize,      \
                                     96 * 1024));                                                                  \
    gather_tile_kernel<TT, SS><<<tgrid, TL_THREADS, smem, s>>>(imgs, (const TT*)feats, lows, tsteps_dev, (TT*)out, \
                                                               B, I1, T, Hp, Wp, phase ? 1 : 0);                       \
  } while (0)
    if (bf) { if (s_int == 1) GT_LAUNCH(bf16, 1); else GT_LAUNCH(bf16, 2); }
    else    { if (s_int == 1) GT_LAUNCH(float, 1); else GT_LAUNCH(float, 2); }
#undef GT_LAUNCH
    VFI_LAUNCHED(h);
    return 0;
  }
#define GATHER_LAUNCH(TT, F)                                                                             \
  gather_kernel<TT, F><<<grid, 128, 0, s>>>(imgs, (const TT*)feats, lows, tsteps_dev, (TT*)out, B, I1, T, Hp, Wp, \
                                            s_int, phase ? 1 : 0)
  if (bf) { if (first) GATHER_LAUNCH(bf16, true); else GATHER_LAUNCH(bf16, false); }
  else    { if (first) GATHER_LAUNCH(float, true); else GATHER_LAUNCH(float, false); }
#undef GATHER_LAUNCH
  VFI_LAUNCHED(h);
  return 0;
}

// ------------------------------------------------------------------------------------------
// IFNet tail + wrapper epilogue, fused (north_star kernel (b)):
//   flow = sum of the five blocks' upsampled deltas, mask = the last block's   (Appendix A.2; state_at)
//   warped0 = warp(img0, flow[:, :2]); warped1 = warp(img1, flow[:, 2:4]); m = sigmoid(mask)
//   merged = warped0*m + warped1*(1-m)
//   unpad (base.py:227-229); (x*255).clip(0,255).astype(uint8) truncation (base.py:188-201)
//   HUD composite frame[mask] = (f0 if t<0.5 else f1)[mask]              (adaptive_pipeline.py:399-409)
// The u8 tile is assembled in shared memory and written as aligned 32-bit words, 96 B per tile row.
// base0 / base1: pixel index of the pair's two frames in the RGBX image set
__device__ __forceinline__ void blend_px(const uint32_t* __restrict__ img32, int base0, int base1, const float4 fl,
                                         const float mk, int x, int y, int Hp, int Wp, float* o /*3*/) {
  const Bil b0 = bil_setup((float)x + fl.x, (float)y + fl.y, Wp, Hp, base0);
  const Bil b1 = bil_setup((float)x + fl.z, (float)y + fl.w, Wp, Hp, base1);
  float a[3], c[3];
  sample_rgbx(img32, b0, a);
  sample_rgbx(img32, b1, c);
  const float m = __fdividef(1.0f, 1.0f + __expf(-mk));  // sigmoid; ~2 ulp, far inside the 1e-4 parity budget
#pragma unroll
  for (int k = 0; k < 3; k++) o[k] = a[k] * m + c[k] * (1.0f - m);
}
__device__ __forceinline__ uint8_t quant_u8(float v) {
  v = fminf(fmaxf(v * 255.0f, 0.f), 255.f);
  return (uint8_t)(int)v;  // truncation, as numpy astype(uint8)
}
// PRE: the flow of all coarser blocks was already summed per pixel by the finest block's gather (flow_pre); `lows` then
// holds only the last block's full-resolution output and nothing is staged.  PRE = 1: that output is fp32 texels
// {flow delta x4, mask, -, -, -}; PRE = 2: the same five values as fp16 (16-byte texels).
template <int PRE>
__global__ void __launch_bounds__(TL_THREADS) final_blend_kernel(const uint8_t* __restrict__ imgs,
                                                                 const __grid_constant__ LowSet lows,
                                                                 const float4* __restrict__ flow_pre,
                                                                 const uint8_t* __restrict__ hud,
                                                                 const float* __restrict__ tsteps,
                                                                 uint8_t* __restrict__ out,
                                                                 float* __restrict__ out_f32,
                                                                 float* __restrict__ flow_out, int B, int I1, int Tn, int H,
                                                                 int W, int Hp, int Wp, int bgr, int Tr) {
  extern __shared__ float4 tl_sm[];
  __shared__ __align__(16) uint8_t s_rgb[TL_H][TL_W * 3];  // the tile's u8 result, stored row-wise as 32-bit words
  // grid (tiles_x * Tr, tiles_y, items / Tr): Tr = T puts the T timesteps of a pair side by side (shared image reads in L2)
  const int bxq = blockIdx.x / Tr;
  const int n = blockIdx.z * Tr + (blockIdx.x - bxq * Tr), tx0 = bxq * TL_W, ty0 = blockIdx.y * TL_H;
  if (!PRE) {
    tile_stage<2>(lows, n, tx0, ty0, Hp, Wp, tl_sm);
    __syncthreads();
  }
  const int b = n / Tn;
  const float t = tsteps[n - b * Tn];
  const size_t plane = (size_t)Hp * Wp;
  const uint8_t* img0 = imgs + (size_t)b * plane * 4;
  const uint8_t* img1 = imgs + (size_t)(I1 + b) * plane * 4;
  const uint8_t* hsrc = t < 0.5f ? img0 : img1;
  // a warp = one tile row (lane = column): conflict-free LDS.128 of the staged rows, broadcast row table
  const int xl = threadIdx.x & (TL_W - 1), x = tx0 + xl;
#pragma unroll 2
  for (int i = 0; i < TL_H / (TL_THREADS / TL_W); i++) {
    const int yl = (threadIdx.x >> 5) + (TL_THREADS / TL_W) * i, y = ty0 + yl;
    if (y >= H || x >= W) continue;
    float o[3];
    float4 m4;
    float4 fl;
    if (PRE) {
      const size_t pix = ((size_t)n * Hp + y) * Wp + x;
      const float4 f3 = __ldg(flow_pre + pix);
      if (PRE == 2) {
        const uint4 tq = __ldg(reinterpret_cast<const uint4*>(lows.p[0]) + pix);
        const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&tq.x));
        const float2 c = __half22float2(*reinterpret_cast<const __half2*>(&tq.y));
        fl = make_float4(a.x, a.y, c.x, c.y);
        m4.x = __low2float(*reinterpret_cast<const __half2*>(&tq.z));
      } else {
        const float4* px = reinterpret_cast<const float4*>(lows.p[0]) + 2 * pix;
        fl = __ldg(px);
        m4.x = __ldg(&px[1].x);
      }
      fl.x += f3.x; fl.y += f3.y; fl.z += f3.z; fl.w += f3.w;
    } else {
      fl = tile_eval<2>(lows, tl_sm, n, x, y, xl, yl, &m4);
    }
    blend_px(reinterpret_cast<const uint32_t*>(imgs), b * (int)plane, (I1 + b) * (int)plane, fl, m4.x, x, y, Hp, Wp, o);
    const size_t opix = ((size_t)n * H + y) * W + x;
    if (out_f32) { out_f32[opix * 3] = o[0]; out_f32[opix * 3 + 1] = o[1]; out_f32[opix * 3 + 2] = o[2]; }
    if (flow_out) reinterpret_cast<float4*>(flow_out)[opix] = fl;
    uint8_t r = quant_u8(o[0]), g = quant_u8(o[1]), bl = quant_u8(o[2]);
    if (hud && hud[((size_t)b * H + y) * W + x]) {
      uchar4 sp = __ldg(reinterpret_cast<const uchar4*>(hsrc) + (size_t)y * Wp + x);
      r = sp.x; g = sp.y; bl = sp.z;
    }
    if (bgr) { const uint8_t sw = r; r = bl; bl = sw; }           // the caller's frames are BGR: uint8 result in its order
    s_rgb[yl][3 * xl] = r; s_rgb[yl][3 * xl + 1] = g; s_rgb[yl][3 * xl + 2] = bl;
  }
  __syncthreads();
  const int cols = min(TL_W, W - tx0), rows = min(TL_H, H - ty0);
  if ((W & 3) == 0) {
    // tx0*3 and W*3 are multiples of 4: each tile row is cols*3/4 aligned words
    const int wpr = cols * 3 / 4;
    for (int w = threadIdx.x; w < TL_H * (TL_W * 3 / 4); w += TL_THREADS) {
      const int row = w / (TL_W * 3 / 4), col = w - row * (TL_W * 3 / 4);
      if (row < rows && col < wpr)
        reinterpret_cast<uint32_t*>(out + (((size_t)n * H + ty0 + row) * W + tx0) * 3)[col] =
            reinterpret_cast<const uint32_t*>(&s_rgb[row][0])[col];
    }
  } else {
    for (int w = threadIdx.x; w < TL_H * TL_W * 3; w += TL_THREADS) {
      const int row = w / (TL_W * 3), col = w - row * (TL_W * 3);
      if (row < rows && col < cols * 3) out[(((size_t)n * H + ty0 + row) * W + tx0) * 3 + col] = s_rgb[row][col];
    }
  }
}
int k_final_blend(vfi_handle* h, const uint8_t* imgs, const LowSet& lows, const float* flow_pre, const uint8_t* hud,
                  const float* tsteps_dev, uint8_t* out, float* out_f32, float* flow_out, int B, int I1, int T, int H, int W, int Hp,
                  int Wp, cudaStream_t s, int bgr) {
  ProfScope _ps(h, PC_BLEND, s);
  const size_t smem = tl_smem_bytes(lows, 2);
  const int Tr = h->opt_t_fast ? T : 1;
  const dim3 grid((W + TL_W - 1) / TL_W * Tr, (H + TL_H - 1) / TL_H, B * T / Tr);
  if (flow_pre) {
    if (lows.n < 1 || lows.s[lows.n - 1] != 1) VFI_FAIL(h, VFI_E_ARG, "final blend: pre-summed flow needs a full-resolution last level");
    LowSet last;
    memset(&last, 0, sizeof(last));
    last.push(lows.p[lows.n - 1], 1, lows.c4[lows.n - 1] * 4, Hp, Wp);
    if (lows.c4[lows.n - 1] == 1)  // fp16 texels
      final_blend_kernel<2><<<grid, TL_THREADS, 0, s>>>(imgs, last, reinterpret_cast<const float4*>(flow_pre), hud, tsteps_dev,
                                                        out, out_f32, flow_out, B, I1, T, H, W, Hp, Wp, bgr, Tr);
    else if (lows.c4[lows.n - 1] == 2)
      final_blend_kernel<1><<<grid, TL_THREADS, 0, s>>>(imgs, last, reinterpret_cast<const float4*>(flow_pre), hud, tsteps_dev,
                                                        out, out_f32, flow_out, B, I1, T, H, W, Hp, Wp, bgr, Tr);
    else VFI_FAIL(h, VFI_E_ARG, "final blend: pre-summed flow needs the compact last level");
  } else {
    final_blend_kernel<0><<<grid, TL_THREADS, smem, s>>>(imgs, lows, nullptr, hud, tsteps_dev, out, out_f32, flow_out, B,
                                                             I1, T, H, W, Hp, Wp, bgr, Tr);
  }
  VFI_LAUNCHED(h);
  return 0;
}

// flow [N,H,W,4] + mask [N,H,W] -> fm [N,Hp,Wp,8] (standalone vfi_warp_blend entry)
__global__ void pack_fm_kernel(const float* __restrict__ flow, const float* __restrict__ mask, float* __restrict__ fm,
                               int N, int H, int W, int Hp, int Wp) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  size_t total = (size_t)N * Hp * Wp;
  if (i >= total) return;
  int x = (int)(i % Wp), y = (int)((i / Wp) % Hp), n = (int)(i / ((size_t)Wp * Hp));
  float4 fl = make_float4(0, 0, 0, 0);
  float mk = 0.f;
  if (x < W && y < H) {
    size_t p = ((size_t)n * H + y) * W + x;
    fl = reinterpret_cast<const float4*>(flow)[p];
    mk = mask[p];
  }
  reinterpret_cast<float4*>(fm)[i * 2] = fl;
  reinterpret_cast<float4*>(fm)[i * 2 + 1] = make_float4(mk, 0, 0, 0);
}
int k_pack_fm(vfi_handle* h, const float* flow, const float* mask, float* fm, int N, int H, int W, int Hp, int Wp,
              cudaStream_t s) {
  ProfScope _ps(h, PC_MISC, s);
  size_t total = (size_t)N * Hp * Wp;
  pack_fm_kernel<<<cdiv(total, 256), 256, 0, s>>>(flow, mask, fm, N, H, W, Hp, Wp);
  VFI_LAUNCHED(h);
  return 0;
}

// ------------------------------------------------------------------------------------------
// test-hook conversions
__global__ void f32_to_bf16_kernel(const float* s, bf16* d, size_t n) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i < n) d[i] = __float2bfloat16(s[i]);
}
__global__ void bf16_to_f32_kernel(const bf16* s, float* d, size_t n) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i < n) d[i] = __bfloat162float(s[i]);
}
int k_f32_to_act(vfi_handle* h, const float* src, void* dst, size_t n, bool bf, cudaStream_t s) {
  if (bf) { f32_to_bf16_kernel<<<cdiv(n, 256), 256, 0, s>>>(src, (bf16*)dst, n); VFI_LAUNCHED(h); }
  else VFI_CUDA(h, cudaMemcpyAsync(dst, src, n * 4, cudaMemcpyDeviceToDevice, s));
  return 0;
}
int k_act_to_f32(vfi_handle* h, const void* src, float* dst, size_t n, bool bf, cudaStream_t s) {
  if (bf) { bf16_to_f32_kernel<<<cdiv(n, 256), 256, 0, s>>>((const bf16*)src, dst, n); VFI_LAUNCHED(h); }
  else VFI_CUDA(h, cudaMemcpyAsync(dst, src, n * 4, cudaMemcpyDeviceToDevice, s));
  return 0;
}
// plain NHWC -> 4 stride-2 phase planes [N,4,H/2,W/2,C]
__global__ void plain_to_phase_kernel(const bf16* __restrict__ s, bf16* __restrict__ d, int N, int H, int W, int C) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  size_t total = (size_t)N * H * W * C;
  if (i >= total) return;
  int c = (int)(i % C);
  size_t p = i / C;
  int x = (int)(p % W), y = (int)((p / W) % H), n = (int)(p / ((size_t)W * H));
  size_t o = ((((size_t)n * 4 + (y & 1) * 2 + (x & 1)) * (H / 2) + (y >> 1)) * (W / 2) + (x >> 1)) * C + c;
  d[o] = s[i];
}
int k_plain_to_phase(vfi_handle* h, const bf16* src, bf16* dst, int N, int H, int W, int C, cudaStream_t s) {
  size_t total = (size_t)N * H * W * C;
  plain_to_phase_kernel<<<cdiv(total, 256), 256, 0, s>>>(src, dst, N, H, W, C);
  VFI_LAUNCHED(h);
  return 0;
}

// ------------------------------------------------------------------------------------------
// Finest block at scale 1.0 (s = 1): block-input gather FUSED with conv0's first stride-2 convolution.
// The 28-channel block input (64 B/pixel in bf16, 1.6 GB per 12 frames at 1080p) never goes to HBM: a CTA gathers
// the 33 x 33 input pixels (32 x 32 tile + the top / left halo a stride-2, pad-1 conv needs) straight into shared
// memory, laid out as conv_s2.cu's four phase-plane windows [Cin/8][plane][17][17][8 ch], one warp issues the
// 2 x 9 x 2 tcgen05.mma (M = 128 = 16 x 8 outputs, N = 16, K = 16) against the 9 KB of weights and four warps run the
// bias + LeakyReLU epilogue, writing the 16 x 16 x 16 output tile as the phase planes conv0's second conv reads.
#define GC_THREADS 384   // 3 pixels per thread cover the 33 x 33 = 1089 input pixels with every warp equally loaded
#define GC_TW 33         // staged / gathered region: the 32 x 32 tile plus its top row and left column
#define GC_PW 17
#define GC_PLANE 292  // plane stride (16-byte units): >= 17 * GC_PW and == 4 mod 8, so the two planes a row's pixels
                      // alternate between sit 64 B apart modulo 128 and a quarter-warp's 16-byte stores never collide
#define GC_NPIX (4 * GC_PLANE)
#define GC_CS 1170  // chunk stride (16-byte units): >= GC_NPIX, == 2 mod 8 (bank spread, see umma::chunk_stride)
#define GC_A_BYTES (4 * GC_CS * 16)
#define GC_W_BYTES (9 * 32 * 16 * 2)
#define GC_TAB (5 * GC_TW)

__device__ __forceinline__ uint32_t gc_pix_index(int ly, int lx) {
  // region-local (ly, lx) in [0, 33): global pixel (ty0 - 1 + ly, tx0 - 1 + lx); ty0, tx0 are even
  const int py = (ly + 1) & 1, px = (lx + 1) & 1;    // parity of the global coordinate
  const int pr = (ly + 1) >> 1, pc = (lx + 1) >> 1;  // window row / col: (g >> 1) - t0/2 + 1
  return (uint32_t)((py * 2 + px) * GC_PLANE + pr * GC_PW + pc);
}
struct GcRows { int start[6]; };  // first staged row of each level; start[n] = all rows
static size_t gc_stage_bytes(const LowSet& lows, GcRows* R, int tw = GC_TW) {
  int rows = 0;
  for (int k = 0; k < 5; k++) {
    R->start[k] = rows;
    if (k < lows.n) rows += tl_rows(lows.s[k]);
  }
  R->start[5] = rows;
  return ((size_t)5 * tw + (size_t)rows * tw) * sizeof(float4);
}
__device__ __forceinline__ Tap ld_tap(const Tap* p) {
  const int4 v = __ldg(reinterpret_cast<const int4*>(p));
  Tap t;
  t.i0 = v.x; t.i1 = v.y; t.h = __int_as_float(v.z); t.l = __int_as_float(v.w);
  return t;
}
// The flow rows of every coarse level for the 33 x 33 region whose origin is (ox, oy) = (tx0 - 1, ty0 - 1), x-interpolated
// to the region's columns and pre-multiplied by the level's flow scale (a power of two: exact).  The staged rows of all
// levels are one flat array of 33-float4 rows, so entry e is simply staged row e / 33, column e % 33; the x taps come
// from the per-call tables (k_build_taps), the level of a row from the host-computed row starts.
template <int TW, int NT, int UNR = 2>
__device__ __forceinline__ void flat_stage(const LowSet& L, const GcRows& R, int n, int ox, int oy, int Hp, int Wp, float4* sm,
                                           int tid = -1) {
  constexpr int GC_TW_ = TW, GC_TAB_ = 5 * TW, GC_THREADS_ = NT;
  if (tid < 0) tid = (int)threadIdx.x;     // NT threads with ids 0 .. NT-1 cooperate (a subset of the CTA's warps may stage)
  RowTap* ytab = reinterpret_cast<RowTap*>(sm);
  const float oyf = (float)max(oy, 0) + 0.5f;
  for (int e = tid; e < L.n * GC_TW_; e += GC_THREADS_) {
    const int k = e / GC_TW_, yl = e - k * GC_TW_;
    const Tap ty = ld_tap(L.yt[k] + min(max(oy + yl, 0), Hp - 1));
    const int r_lo = (int)fmaxf(L.rs[k] * oyf - 0.5f, 0.f);
    const int off = GC_TAB_ + R.start[k] * GC_TW_;
    RowTap t;
    t.o0 = (off + (ty.i0 - r_lo) * GC_TW_) * 16; t.o1 = (off + (ty.i1 - r_lo) * GC_TW_) * 16;  // byte offsets
    t.hy = ty.h; t.ly = ty.l;
    ytab[e] = t;
  }
  const int total = R.start[5] * GC_TW_;
#pragma unroll UNR   // loads in flight per thread: 4 takes 2.8 % off the persistent fused finest gather (8: 8 % slower); 2 elsewhere
  for (int e = tid; e < total; e += GC_THREADS_) {
    const int u = e / GC_TW_, col = e - u * GC_TW_;
    int k = 0;
#pragma unroll
    for (int j = 1; j < 5; j++) k += (j < L.n && u >= R.start[j]) ? 1 : 0;
    const int Hl = L.Hl[k], Wl = L.Wl[k], c4 = L.c4[k];
    const int r_lo = (int)fmaxf(L.rs[k] * oyf - 0.5f, 0.f);
    const int ry = min(r_lo + u - R.start[k], Hl - 1);
    const Tap tx = ld_tap(L.xt[k] + min(max(ox + col, 0), Wp - 1));
    const float4* row = reinterpret_cast<const float4*>(L.p[k]) + ((size_t)n * Hl + ry) * Wl * c4;
    const float4 a = __ldg(row + tx.i0 * c4), b = __ldg(row + tx.i1 * c4);
    const float fs = L.fs[k];
    sm[GC_TAB_ + e] = make_float4((tx.h * a.x + tx.l * b.x) * fs, (tx.h * a.y + tx.l * b.y) * fs,
                                 (tx.h * a.z + tx.l * b.z) * fs, (tx.h * a.w + tx.l * b.w) * fs);
  }
}
// flow at region-local pixel (lx, ly) from the staged rows (every level of a finest block's LowSet has s > 1)
template <int TW>
__device__ __forceinline__ float4 flat_eval(const LowSet& L, const float4* sm, int lx, int ly) {
  const RowTap* ytab = reinterpret_cast<const RowTap*>(sm);
  float4 fl = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
  for (int k = 0; k < 5; k++) {
    if (k < L.n) {
      const RowTap t = ytab[k * TW + ly];
      const char* col = reinterpret_cast<const char*>(sm + lx);
      const float4 a = *reinterpret_cast<const float4*>(col + t.o0), c = *reinterpret_cast<const float4*>(col + t.o1);
      fl.x = fmaf(t.ly, c.x, fmaf(t.hy, a.x, fl.x)); fl.y = fmaf(t.ly, c.y, fmaf(t.hy, a.y, fl.y));
      fl.z = fmaf(t.ly, c.z, fmaf(t.hy, a.z, fl.z)); fl.w = fmaf(t.ly, c.w, fmaf(t.hy, a.w, fl.w));
    }
  }
  return fl;
}

template <typename T>
__global__ void __launch_bounds__(GC_THREADS, 2)
gather_conv0a_kernel(const uint8_t* __restrict__ imgs, const T* __restrict__ feats, const __grid_constant__ LowSet lows,
                     const float* __restrict__ tsteps, const bf16* __restrict__ wpack, const float* __restrict__ bias,
                     bf16* __restrict__ out, float4* __restrict__ flow_sum, const __grid_constant__ GcRows grows, int B, int I1, int Tn, int Hp,
                     int Wp, float slope, int stage_bytes, int ablate) {
  using namespace umma;
  extern __shared__ float4 tl_sm[];  // staged flow rows | A tile | weights
  __shared__ uint64_t mma_bar;
  __shared__ uint32_t tmem_slot;
  __shared__ Tap mtab[2][GC_TW];  // x / y taps of the mask / feat level for the region's columns / rows
  uint8_t* a_s = reinterpret_cast<uint8_t*>(tl_sm) + stage_bytes;
  const int n = blockIdx.z, tx0 = blockIdx.x * TL_W, ty0 = blockIdx.y * TL_H;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint8_t* w_s = a_s + GC_A_BYTES;
  if (threadIdx.x == 0) { bar_init(&mma_bar, 1); fence_barrier_init(); }
  if (warp == 0) tmem_alloc(&tmem_slot, 32);
  for (int i = threadIdx.x; i < GC_W_BYTES / 16; i += GC_THREADS)
    reinterpret_cast<uint4*>(w_s)[i] = __ldg(reinterpret_cast<const uint4*>(wpack) + i);
  // stage only the flow rows (incl. the halo row / column); mask / feat of the finest coarse level are interpolated
  // straight from global memory (L1 / L2 resident: one 64 B texel serves four pixels), which leaves room for the A tile
  // beside the staging area at two CTAs per SM and lets a pixel go from evaluation to its A chunks in one pass
  if (!(ablate & 1)) flat_stage<GC_TW, GC_THREADS>(lows, grows, n, tx0 - 1, ty0 - 1, Hp, Wp, tl_sm);
  if (threadIdx.x >= GC_THREADS - 2 * GC_TW) {  // the last warps: they have the fewest staging entries
    const int e = threadIdx.x - (GC_THREADS - 2 * GC_TW), ax = e / GC_TW, j = e - ax * GC_TW;
    const int kl = lows.n - 1;
    const Tap* src = ax ? lows.yt[kl] + max(ty0 - 1 + j, 0) : lows.xt[kl] + max(tx0 - 1 + j, 0);
    reinterpret_cast<int4*>(&mtab[ax][j])[0] = __ldg(reinterpret_cast<const int4*>(src));
  }
  tc_before();
  __syncthreads();
  tc_after();
  const uint32_t tmem_base = tmem_slot;

  const int b = n / Tn;
  const float t = tsteps[n - b * Tn];
  const int plane = Hp * Wp;
  const uint32_t* img32 = reinterpret_cast<const uint32_t*>(imgs);
  const int base0 = b * plane, base1 = (I1 + b) * plane;
  const int kl = lows.n - 1;  // finest coarse level: mask / feat come from it (16-channel texels: c4 == 4)
  const int mWl = lows.Wl[kl];
  const bf16 *mfp = lows.mf[kl], *f7p = lows.f7[kl];
  const int mtb = n * lows.Hl[kl] * mWl;
  float4* fsum = flow_sum ? flow_sum + (size_t)n * plane : nullptr;
  // Work units: region rows 0..32 with lane = column 1..32 (one image row per warp instruction: two or three cache
  // lines per tap instead of the five a row-straddling warp touches), then the halo column (lx = 0) in two more units.
  constexpr int GC_UNITS = GC_TW + 2;
#pragma unroll 1
  for (int unit = warp; unit < GC_UNITS; unit += GC_THREADS / 32) {
    int ly, lx;
    if (unit < GC_TW) { ly = unit; lx = lane + 1; }
    else { ly = (unit - GC_TW) * 32 + lane; lx = 0; if (ly >= GC_TW) continue; }
    const int gy = ty0 - 1 + ly, gx = tx0 - 1 + lx;  // global pixel
    uint4 u[4];
#pragma unroll
    for (int j = 0; j < 4; j++) u[j] = make_uint4(0u, 0u, 0u, 0u);
    if (gy >= 0 && gx >= 0 && !(ablate & 2)) {  // else: outside the image = the convolution's zero padding
      const float4 fl = (ablate & 8) ? make_float4(0.5f, 0.25f, -0.5f, -0.25f) : flat_eval<GC_TW>(lows, tl_sm, lx, ly);
      // the flow of blocks 0..3 at this pixel, for the final blend (tile-interior pixels: each written exactly once)
      if (fsum && ly > 0 && lx > 0) fsum[gy * Wp + gx] = fl;
      // mask + 8 feature channels: bilinear x2 of texel channels 4..12
      const Tap ty = mtab[1][ly], tx = mtab[0][lx];
      const int r0 = mtb + ty.i0 * mWl, r1 = mtb + ty.i1 * mWl;
      const float w00 = ty.h * tx.h, w01 = ty.h * tx.l, w10 = ty.l * tx.h, w11 = ty.l * tx.l;
      float mf[9];
      {
        float a[9], c[9];
        mf9_load(mfp, f7p, r0 + tx.i0, a);
        mf9_load(mfp, f7p, r0 + tx.i1, c);
#pragma unroll
        for (int i = 0; i < 9; i++) mf[i] = w00 * a[i] + w01 * c[i];
        mf9_load(mfp, f7p, r1 + tx.i0, a);
        mf9_load(mfp, f7p, r1 + tx.i1, c);
#pragma unroll
        for (int i = 0; i < 9; i++) mf[i] += w10 * a[i] + w11 * c[i];
      }
      const Bil b0 = bil_setup((float)gx + fl.x, (float)gy + fl.y, Wp, Hp, base0);
      const Bil b1 = bil_setup((float)gx + fl.z, (float)gy + fl.w, Wp, Hp, base1);
      float c0[3], c1[3], v0[4], v1[4];
      sample_rgbx(img32, b0, c0);
      sample_rgbx(img32, b1, c1);
      sample_feat_px<T>(feats, b0, v0);
      sample_feat_px<T>(feats, b1, v1);
      // channel order of the block input: warped img0 (3), img1 (3), f0 (4), f1 (4), timestep, mask, feat (8), flow (4)
      u[0] = make_uint4(pk_bf16(c0[0], c0[1]), pk_bf16(c0[2], c1[0]), pk_bf16(c1[1], c1[2]), pk_bf16(v0[0], v0[1]));
      u[1] = make_uint4(pk_bf16(v0[2], v0[3]), pk_bf16(v1[0], v1[1]), pk_bf16(v1[2], v1[3]), pk_bf16(t, mf[0]));
      u[2] = make_uint4(pk_bf16(mf[1], mf[2]), pk_bf16(mf[3], mf[4]), pk_bf16(mf[5], mf[6]), pk_bf16(mf[7], mf[8]));
      u[3] = make_uint4(pk_bf16(fl.x, fl.y), pk_bf16(fl.z, fl.w), 0u, 0u);
    }
    const uint32_t pix = gc_pix_index(ly, lx);
#pragma unroll
    for (int j = 0; j < 4; j++) *reinterpret_cast<uint4*>(a_s + ((size_t)j * GC_CS + pix) * 16) = u[j];
  }
  fence_async_smem();
  __syncthreads();

  // ---- phase 3: 2 sub-tiles (output columns 0-7, 8-15) x 9 taps x 2 k-steps of tcgen05.mma, then the epilogue ----
  if (ablate & 4) {
    if (warp == 0) { tc_after(); tmem_dealloc(tmem_base, 32); }
    return;
  }
  if (warp == 0) {
    const uint32_t leader = elect_one();
    constexpr uint32_t IDESC = idesc_bf16_m128(16);
    const uint64_t ad0 = desc_ns(s_u32(a_s), GC_CS * 16, GC_PW * 16);
    const uint64_t bd0 = desc_ns(s_u32(w_s), 16 * 16, 128);
#pragma unroll
    for (int tap = 0; tap < 9; tap++) {
#pragma unroll
      for (int k = 0; k < 2; k++) {
#pragma unroll
        for (int j = 0; j < 2; j++) {
          const int ky = tap / 3, kx = tap % 3;
          const int pl = (ky != 1 ? 2 : 0) + (kx != 1 ? 1 : 0);
          const uint32_t a_off = (uint32_t)(pl * GC_PLANE + (ky != 0 ? 1 : 0) * GC_PW + (kx != 0 ? 1 : 0) + 8 * j + 2 * k * GC_CS);
          const uint32_t b_off = (uint32_t)((tap * 4 + 2 * k) * 16);
          mma_bf16_elect(leader, tmem_base + (uint32_t)(j * 16), ad0 + a_off, bd0 + b_off, IDESC, (tap | k) ? 1u : 0u);
        }
      }
    }
    if (leader) commit_to(&mma_bar);
    __syncwarp();
  }
  if (warp < 4) {
    bar_wait(&mma_bar, 0);
    tc_after();
    const int m = warp * 32 + lane;
    const int Ho = Hp >> 1, Wo = Wp >> 1;
    const int oy = (ty0 >> 1) + (m >> 3);
    float v[2][16];
    ld16(tmem_base + ((uint32_t)(warp * 32) << 16), v[0]);
    ld16(tmem_base + ((uint32_t)(warp * 32) << 16) + 16, v[1]);
#pragma unroll
    for (int j = 0; j < 2; j++) {
      const int ox = (tx0 >> 1) + 8 * j + (m & 7);
#pragma unroll
      for (int c = 0; c < 16; c++) {
        float x = v[j][c] + __ldg(bias + c);
        v[j][c] = x > 0.f ? x : x * slope;
      }
      // conv0's second conv is stride 2 too: write its four phase planes [N, 4, Ho/2, Wo/2, 16]
      const size_t opix = ((((size_t)n * 4 + (oy & 1) * 2 + (ox & 1)) * (Ho >> 1) + (oy >> 1)) * (Wo >> 1) + (ox >> 1));
      uint4 o0, o1;
      o0.x = pk_bf16(v[j][0], v[j][1]); o0.y = pk_bf16(v[j][2], v[j][3]); o0.z = pk_bf16(v[j][4], v[j][5]); o0.w = pk_bf16(v[j][6], v[j][7]);
      o1.x = pk_bf16(v[j][8], v[j][9]); o1.y = pk_bf16(v[j][10], v[j][11]); o1.z = pk_bf16(v[j][12], v[j][13]); o1.w = pk_bf16(v[j][14], v[j][15]);
      st_v8(out + opix * 16, o0, o1);
    }
  }
  tc_before();
  __syncthreads();
  if (warp == 0) {
    tc_after();
    tmem_dealloc(tmem_base, 32);
  }
}

// Persistent form of the kernel above (2 CTAs per SM loop over the tiles): barriers, TMEM and the 9 KB of weights are set
// up once per CTA, and while warps 0-3 issue tile i's MMAs and run its epilogue, warps 4-11 already stage the flow rows of
// tile i+1 -- the two latency-bound phases overlap instead of adding (phase ablation on B200, 24 items of 1080p: set-up
// 0.37 ms + staging 0.41 + gather 1.22 + MMA / epilogue 0.33 = the one-tile kernel's 2.28 ms).
// ORDER: the tile order, a compile-time choice (the kernel sits at its 80-register cap; run-time digit roles cost spills)
template <typename T, int ORDER>
__global__ void __launch_bounds__(GC_THREADS, 2)
gather_conv0a_p_kernel(const uint8_t* __restrict__ imgs, const T* __restrict__ feats, const __grid_constant__ LowSet lows,
                       const float* __restrict__ tsteps, const bf16* __restrict__ wpack, const float* __restrict__ bias,
                       bf16* __restrict__ out, float4* __restrict__ flow_sum, const __grid_constant__ GcRows grows, int B, int I1, int Tn, int Hp,
                       int Wp, float slope, int stage_bytes, int tiles_x, int tiles_y, int num_tiles) {
  constexpr int tfast = ORDER;
  using namespace umma;
  extern __shared__ float4 tl_sm[];  // staged flow rows | A tile | weights
  __shared__ uint64_t mma_bar;
  __shared__ uint32_t tmem_slot;
  __shared__ Tap mtab[2][GC_TW];
  uint8_t* a_s = reinterpret_cast<uint8_t*>(tl_sm) + stage_bytes;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint8_t* w_s = a_s + GC_A_BYTES;
  constexpr int ablate = 0;
  (void)ablate;
  if (threadIdx.x == 0) { bar_init(&mma_bar, 1); fence_barrier_init(); }
  if (warp == 0) tmem_alloc(&tmem_slot, 32);
  for (int i = threadIdx.x; i < GC_W_BYTES / 16; i += GC_THREADS)
    reinterpret_cast<uint4*>(w_s)[i] = __ldg(reinterpret_cast<const uint4*>(wpack) + i);
  const int kl = lows.n - 1;  // finest coarse level: mask / feat come from it (16-channel texels: c4 == 4)
  int tile = blockIdx.x;
  // stage tile `t` (flow rows + the mask / feat tap tables) with the NT threads whose ids are tid = 0 .. NT-1
  auto stage = [&](int tx, int ty, int nn, auto nt_tag, int tid) {
    constexpr int NT = decltype(nt_tag)::value;
    flat_stage<GC_TW, NT, 4>(lows, grows, nn, tx * TL_W - 1, ty * TL_H - 1, Hp, Wp, tl_sm, tid);
    if (tid >= NT - 2 * GC_TW) {
      const int e = tid - (NT - 2 * GC_TW), ax = e / GC_TW, j = e - ax * GC_TW;
      const Tap* src = ax ? lows.yt[kl] + max(ty * TL_H - 1 + j, 0) : lows.xt[kl] + max(tx * TL_W - 1 + j, 0);
      reinterpret_cast<int4*>(&mtab[ax][j])[0] = __ldg(reinterpret_cast<const int4*>(src));
    }
  };
  // tile -> (tx, ty, n) once; afterwards the mixed-radix coordinates advance by the (decomposed) grid stride: no integer
  // divisions in the loop (they were 6 % of the kernel's instructions)
  // Tile order (tfast): TIMESTEP fastest, then x, y, pair -- the T items of a pair warp the same two images (features and
  // RGBX, 24 B per pixel and image), so the CTAs working side by side read one image region for all T timesteps while it is
  // in L2; item-major order re-read every image from DRAM once per timestep (one item's reads + writes ~ the L2 size).
  // tfast = 0 keeps the item-major order: the timestep digit then has radix 1 and the top digit is the item.
  // tfast = 2: (timestep, pair, x, y) -- in clip mode pair b's second frame is pair b + 1's first, so with the pairs side by
  // side too each frame is read once instead of twice.
  const int Tr = tfast ? Tn : 1;
  const int R1 = tfast == 2 ? B : tiles_x, R2 = tfast == 2 ? tiles_x : tiles_y;   // radices of the digits after the timestep
  int ct = tile % Tr, d1 = (tile / Tr) % R1, d2 = (tile / (Tr * R1)) % R2, d3 = tile / (Tr * R1 * R2);
  const int g = (int)gridDim.x, gst = g % Tr, gs1 = (g / Tr) % R1, gs2 = (g / (Tr * R1)) % R2, gs3 = g / (Tr * R1 * R2);
  int ctx = tfast == 2 ? d2 : d1, cty = tfast == 2 ? d3 : d2, ctop = tfast == 2 ? d1 : d3;
  int cn = ctop * Tr + ct;                                 // item = pair * T + timestep
  if (tile < num_tiles) stage(ctx, cty, cn, std::integral_constant<int, GC_THREADS>{}, (int)threadIdx.x);
  fence_async_smem();
  tc_before();
  __syncthreads();
  tc_after();
  const uint32_t tmem_base = tmem_slot;
  const int plane = Hp * Wp;
  const uint32_t* img32 = reinterpret_cast<const uint32_t*>(imgs);
  const int mWl = lows.Wl[kl];
  const bf16 *mfp = lows.mf[kl], *f7p = lows.f7[kl];
  for (int it = 0; tile < num_tiles; tile += g, it++) {
    const int tx0 = ctx * TL_W, ty0 = cty * TL_H, n = cn;
    const int b = tfast ? ctop : n / Tn;
    const float t = tsteps[n - b * Tn];
    // the next tile of this CTA
    ct += gst; d1 += gs1; d2 += gs2; d3 += gs3;
    if (ct >= Tr) { ct -= Tr; d1++; }
    if (d1 >= R1) { d1 -= R1; d2++; }
    if (d2 >= R2) { d2 -= R2; d3++; }
    ctx = tfast == 2 ? d2 : d1; cty = tfast == 2 ? d3 : d2; ctop = tfast == 2 ? d1 : d3;
    cn = ctop * Tr + ct;
    const int base0 = b * plane, base1 = (I1 + b) * plane;
    const int mtb = n * lows.Hl[kl] * mWl;
    float4* fsum = flow_sum ? flow_sum + (size_t)n * plane : nullptr;
    // Work units: region rows 0..32 with lane = column 1..32 (one image row per warp instruction: two or three cache
    // lines per tap instead of the five a row-straddling warp touches), then the halo column (lx = 0) in two more units.
    constexpr int GC_UNITS = GC_TW + 2;
  #pragma unroll 1
    for (int unit = warp; unit < GC_UNITS; unit += GC_THREADS / 32) {
      int ly, lx;
      if (unit < GC_TW) { ly = unit; lx = lane + 1; }
      else { ly = (unit - GC_TW) * 32 + lane; lx = 0; if (ly >= GC_TW) continue; }
      const int gy = ty0 - 1 + ly, gx = tx0 - 1 + lx;  // global pixel
      uint4 u[4];
  #pragma unroll
      for (int j = 0; j < 4; j++) u[j] = make_uint4(0u, 0u, 0u, 0u);
      if (gy >= 0 && gx >= 0) {  // else: outside the image = the convolution's zero padding
        const float4 fl = flat_eval<GC_TW>(lows, tl_sm, lx, ly);
        // the flow of blocks 0..3 at this pixel, for the final blend (tile-interior pixels: each written exactly once)
        if (fsum && ly > 0 && lx > 0) fsum[gy * Wp + gx] = fl;
        // mask + 8 feature channels: bilinear x2 of texel channels 4..12
        const Tap ty = mtab[1][ly], tx = mtab[0][lx];
        const int r0 = mtb + ty.i0 * mWl, r1 = mtb + ty.i1 * mWl;
        const float w00 = ty.h * tx.h, w01 = ty.h * tx.l, w10 = ty.l * tx.h, w11 = ty.l * tx.l;
        float mf[9];
        {
          float a[9], c[9];
          mf9_load(mfp, f7p, r0 + tx.i0, a);
          mf9_load(mfp, f7p, r0 + tx.i1, c);
  #pragma unroll
          for (int i = 0; i < 9; i++) mf[i] = w00 * a[i] + w01 * c[i];
          mf9_load(mfp, f7p, r1 + tx.i0, a);
          mf9_load(mfp, f7p, r1 + tx.i1, c);
  #pragma unroll
          for (int i = 0; i < 9; i++) mf[i] += w10 * a[i] + w11 * c[i];
        }
        const Bil b0 = bil_setup((float)gx + fl.x, (float)gy + fl.y, Wp, Hp, base0);
        const Bil b1 = bil_setup((float)gx + fl.z, (float)gy + fl.w, Wp, Hp, base1);
        float c0[3], c1[3], v0[4], v1[4];
        sample_rgbx(img32, b0, c0);
        sample_rgbx(img32, b1, c1);
        sample_feat_px<T>(feats, b0, v0);
        sample_feat_px<T>(feats, b1, v1);
        // channel order of the block input: warped img0 (3), img1 (3), f0 (4), f1 (4), timestep, mask, feat (8), flow (4)
        u[0] = make_uint4(pk_bf16(c0[0], c0[1]), pk_bf16(c0[2], c1[0]), pk_bf16(c1[1], c1[2]), pk_bf16(v0[0], v0[1]));
        u[1] = make_uint4(pk_bf16(v0[2], v0[3]), pk_bf16(v1[0], v1[1]), pk_bf16(v1[2], v1[3]), pk_bf16(t, mf[0]));
        u[2] = make_uint4(pk_bf16(mf[1], mf[2]), pk_bf16(mf[3], mf[4]), pk_bf16(mf[5], mf[6]), pk_bf16(mf[7], mf[8]));
        u[3] = make_uint4(pk_bf16(fl.x, fl.y), pk_bf16(fl.z, fl.w), 0u, 0u);
      }
      const uint32_t pix = gc_pix_index(ly, lx);
  #pragma unroll
      for (int j = 0; j < 4; j++) *reinterpret_cast<uint4*>(a_s + ((size_t)j * GC_CS + pix) * 16) = u[j];
    }
    fence_async_smem();
    __syncthreads();               // the A tile is complete; the staging area and the tap tables are free again
    if (warp == 0) {
      const uint32_t leader = elect_one();
      constexpr uint32_t IDESC = idesc_bf16_m128(16);
      const uint64_t ad0 = desc_ns(s_u32(a_s), GC_CS * 16, GC_PW * 16);
      const uint64_t bd0 = desc_ns(s_u32(w_s), 16 * 16, 128);
  #pragma unroll
      for (int tap = 0; tap < 9; tap++) {
  #pragma unroll
        for (int k = 0; k < 2; k++) {
  #pragma unroll
          for (int j = 0; j < 2; j++) {
            const int ky = tap / 3, kx = tap % 3;
            const int pl = (ky != 1 ? 2 : 0) + (kx != 1 ? 1 : 0);
            const uint32_t a_off = (uint32_t)(pl * GC_PLANE + (ky != 0 ? 1 : 0) * GC_PW + (kx != 0 ? 1 : 0) + 8 * j + 2 * k * GC_CS);
            const uint32_t b_off = (uint32_t)((tap * 4 + 2 * k) * 16);
            mma_bf16_elect(leader, tmem_base + (uint32_t)(j * 16), ad0 + a_off, bd0 + b_off, IDESC, (tap | k) ? 1u : 0u);
          }
        }
      }
      if (leader) commit_to(&mma_bar);
      __syncwarp();
    }
    if (warp < 4) {
      bar_wait(&mma_bar, (uint32_t)(it & 1));
      tc_after();
      const int m = warp * 32 + lane;
      const int Ho = Hp >> 1, Wo = Wp >> 1;
      const int oy = (ty0 >> 1) + (m >> 3);
      float v[2][16];
      ld16(tmem_base + ((uint32_t)(warp * 32) << 16), v[0]);
      ld16(tmem_base + ((uint32_t)(warp * 32) << 16) + 16, v[1]);
  #pragma unroll
      for (int j = 0; j < 2; j++) {
        const int ox = (tx0 >> 1) + 8 * j + (m & 7);
  #pragma unroll
        for (int c = 0; c < 16; c++) {
          float x = v[j][c] + __ldg(bias + c);
          v[j][c] = x > 0.f ? x : x * slope;
        }
        // conv0's second conv is stride 2 too: write its four phase planes [N, 4, Ho/2, Wo/2, 16]
        const size_t opix = ((((size_t)n * 4 + (oy & 1) * 2 + (ox & 1)) * (Ho >> 1) + (oy >> 1)) * (Wo >> 1) + (ox >> 1));
        uint4 o0, o1;
        o0.x = pk_bf16(v[j][0], v[j][1]); o0.y = pk_bf16(v[j][2], v[j][3]); o0.z = pk_bf16(v[j][4], v[j][5]); o0.w = pk_bf16(v[j][6], v[j][7]);
        o1.x = pk_bf16(v[j][8], v[j][9]); o1.y = pk_bf16(v[j][10], v[j][11]); o1.z = pk_bf16(v[j][12], v[j][13]); o1.w = pk_bf16(v[j][14], v[j][15]);
        st_v8(out + opix * 16, o0, o1);
      }
    }
    if (warp >= 4 && tile + (int)gridDim.x < num_tiles)
      stage(ctx, cty, cn, std::integral_constant<int, GC_THREADS - 128>{}, (int)threadIdx.x - 128);
    tc_before();
    __syncthreads();               // next tile staged; this tile's MMAs complete (the epilogue waited for them), TMEM read out
    tc_after();
  }
  if (warp == 0) tmem_dealloc(tmem_base, 32);
}

// Lb0 = the block's conv0[0] layer (28 -> 16, stride 2).  Output: its activation as the phase planes of [N, Hp/2, Wp/2, 16].
int k_gather_conv0a(vfi_handle* h, const Layer& Lb0, const uint8_t* imgs, const void* feats, const LowSet& lows,
                    const float* tsteps_dev, bf16* out, float* flow_sum, int B, int I1, int T, int Hp, int Wp,
                    cudaStream_t s) {
  ProfScope _ps(h, PC_GATHER_CONV, s);
  if (!Lb0.ws2 || Lb0.cin_pad != 32 || Lb0.cout != 16) VFI_FAIL(h, VFI_E_STATE, "gather_conv0a: layer not packed");
  if (Hp % TL_H || Wp % TL_W) VFI_FAIL(h, VFI_E_ARG, "gather_conv0a: padded size must be a multiple of 32");
  for (int k = 0; k < lows.n; k++)
    if (lows.s[k] < 2) VFI_FAIL(h, VFI_E_ARG, "gather_conv0a: coarse levels only");
  if (lows.n < 1 || !lows.mf[lows.n - 1] || !lows.xt[0]) VFI_FAIL(h, VFI_E_ARG, "gather_conv0a: needs tap tables and plane-stored coarse levels");
  GcRows grows;
  const size_t un = (gc_stage_bytes(lows, &grows) + 127) & ~(size_t)127;
  const size_t smem = un + GC_A_BYTES + GC_W_BYTES;
  VFI_CUDA(h, cudaFuncSetAttribute(gather_conv0a_kernel<bf16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024));
  if (smem > 112 * 1024) VFI_FAIL(h, VFI_E_ARG, "gather_conv0a: shared memory budget");
  if (h->opt_gc_persistent && !h->opt_gc_ablate) {
    const int tiles_x = Wp / TL_W, tiles_y = Hp / TL_H, num_tiles = tiles_x * tiles_y * B * T;
#define GC_P_LAUNCH(ORDER)                                                                                                        \
  do {                                                                                                                            \
    VFI_CUDA(h, cudaFuncSetAttribute(gather_conv0a_p_kernel<bf16, ORDER>, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024)); \
    gather_conv0a_p_kernel<bf16, ORDER><<<std::min(num_tiles, 2 * h->sm_count), GC_THREADS, smem, s>>>(                          \
        imgs, (const bf16*)feats, lows, tsteps_dev, Lb0.ws2, Lb0.bias, out, reinterpret_cast<float4*>(flow_sum), grows, B, I1, T, \
        Hp, Wp, h->slope, (int)un, tiles_x, tiles_y, num_tiles);                                                                  \
  } while (0)
    if (h->opt_t_fast == 0) GC_P_LAUNCH(0);
    else if (h->opt_t_fast == 1) GC_P_LAUNCH(1);
    else GC_P_LAUNCH(2);
#undef GC_P_LAUNCH
    VFI_LAUNCHED(h);
    return 0;
  }
  const dim3 grid(Wp / TL_W, Hp / TL_H, B * T);
  gather_conv0a_kernel<bf16><<<grid, GC_THREADS, smem, s>>>(imgs, (const bf16*)feats, lows, tsteps_dev, Lb0.ws2, Lb0.bias, out,
                                                          reinterpret_cast<float4*>(flow_sum), grows, B, I1, T, Hp, Wp, h->slope, (int)un,
                                                          h->opt_gc_ablate);
  VFI_LAUNCHED(h);
  return 0;
}

// ------------------------------------------------------------------------------------------
// Block-input gather at s = 2 (bf16 mode: the coarse levels are LowPlanes).  Same result as gather_tile_kernel<bf16, 2>,
// restructured around the L1 / issue limits its profile showed:
//   * the flow rows are staged by flat_stage (per-call tap tables, levels in one flat array, pre-scaled);
//   * mask / feat are linear in the coarse texels, so the mean over the 2 x 2 full-resolution positions is ONE bilinear
//     fetch with the mean weights (both positions of a pair share their source cell for every level scale >= 4) taken
//     straight from the bf16 planes -- they are neither staged nor evaluated four times;
//   * 32-bit pixel indexing (one IMAD.WIDE per tap), 80 registers: three CTAs per SM.
#define G2_THREADS 256
__global__ void __launch_bounds__(G2_THREADS, 3)
gather_s2_kernel(const uint8_t* __restrict__ imgs, const bf16* __restrict__ feats, const __grid_constant__ LowSet lows,
                 const __grid_constant__ GcRows grows, const float* __restrict__ tsteps, bf16* __restrict__ out, int B, int I1,
                 int Tn, int Hp, int Wp, int phase, int Tr) {
  extern __shared__ float4 tl_sm[];
  // grid (tiles_x * Tr, tiles_y, items / Tr): with Tr = T the T timesteps of a pair sit side by side in launch order and share
  // the pair's image reads in L2 (see gather_conv0a_p_kernel); Tr = 1 is the item-major order
  const int bxq = blockIdx.x / Tr;
  const int n = blockIdx.z * Tr + (blockIdx.x - bxq * Tr), tx0 = bxq * TL_W, ty0 = blockIdx.y * TL_H;
  flat_stage<TL_W, G2_THREADS>(lows, grows, n, tx0, ty0, Hp, Wp, tl_sm);
  __syncthreads();
  const int b = n / Tn;
  const float t = tsteps[n - b * Tn];
  const int plane = Hp * Wp;
  const uint32_t* img32 = reinterpret_cast<const uint32_t*>(imgs);
  const int base0 = b * plane, base1 = (I1 + b) * plane;
  const int Xl = threadIdx.x & (TL_W / 2 - 1), Yl = threadIdx.x >> 4;
  const int X = (tx0 >> 1) + Xl, Y = (ty0 >> 1) + Yl;
  const int gx = tx0 + 2 * Xl, gy = ty0 + 2 * Yl;
  float mf[9];
  {
    const int kl = lows.n - 1, Wl = lows.Wl[kl];
    const Tap xa = ld_tap(lows.xt[kl] + gx), xb = ld_tap(lows.xt[kl] + gx + 1);
    const Tap ya = ld_tap(lows.yt[kl] + gy), yb = ld_tap(lows.yt[kl] + gy + 1);
    const float hx = 0.5f * (xa.h + xb.h), lx = 0.5f * (xa.l + xb.l), hy = 0.5f * (ya.h + yb.h), ly = 0.5f * (ya.l + yb.l);
    const int r0 = (n * lows.Hl[kl] + ya.i0) * Wl, r1 = (n * lows.Hl[kl] + ya.i1) * Wl;
    const float w00 = hy * hx, w01 = hy * lx, w10 = ly * hx, w11 = ly * lx;
    float a[9], c[9];
    mf9_load(lows.mf[kl], lows.f7[kl], r0 + xa.i0, a);
    mf9_load(lows.mf[kl], lows.f7[kl], r0 + xa.i1, c);
#pragma unroll
    for (int i = 0; i < 9; i++) mf[i] = w00 * a[i] + w01 * c[i];
    mf9_load(lows.mf[kl], lows.f7[kl], r1 + xa.i0, a);
    mf9_load(lows.mf[kl], lows.f7[kl], r1 + xa.i1, c);
#pragma unroll
    for (int i = 0; i < 9; i++) mf[i] += w10 * a[i] + w11 * c[i];
  }
  float c0[3] = {0.f, 0.f, 0.f}, c1[3] = {0.f, 0.f, 0.f}, v0[4] = {0.f, 0.f, 0.f, 0.f}, v1[4] = {0.f, 0.f, 0.f, 0.f};
  float4 fs = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 1
  for (int d = 0; d < 4; d++) {
    // odd groups of four outputs visit their right-hand tap first: the 8 lanes of an LDS.128 phase then hit 8 distinct
    // 16-byte bank groups
    const int xl = 2 * Xl + ((d & 1) ^ ((Xl >> 2) & 1)), yl = 2 * Yl + (d >> 1);
    const float4 fl = flat_eval<TL_W>(lows, tl_sm, xl, yl);
    const Bil b0 = bil_setup((float)(tx0 + xl) + fl.x, (float)(ty0 + yl) + fl.y, Wp, Hp, base0);
    const Bil b1 = bil_setup((float)(tx0 + xl) + fl.z, (float)(ty0 + yl) + fl.w, Wp, Hp, base1);
    float r[4];
    sample_rgbx(img32, b0, r);
    c0[0] += r[0]; c0[1] += r[1]; c0[2] += r[2];
    sample_rgbx(img32, b1, r);
    c1[0] += r[0]; c1[1] += r[1]; c1[2] += r[2];
    sample_feat_px<bf16>(feats, b0, r);
    v0[0] += r[0]; v0[1] += r[1]; v0[2] += r[2]; v0[3] += r[3];
    sample_feat_px<bf16>(feats, b1, r);
    v1[0] += r[0]; v1[1] += r[1]; v1[2] += r[2]; v1[3] += r[3];
    fs.x += fl.x; fs.y += fl.y; fs.z += fl.z; fs.w += fl.w;
  }
  const float q = 0.25f, qf = 0.125f;  // mean of the four positions; the flow also scales by 1/s = 1/2
  uint4 u[4];
  u[0] = make_uint4(umma::pk_bf16(c0[0] * q, c0[1] * q), umma::pk_bf16(c0[2] * q, c1[0] * q), umma::pk_bf16(c1[1] * q, c1[2] * q),
                    umma::pk_bf16(v0[0] * q, v0[1] * q));
  u[1] = make_uint4(umma::pk_bf16(v0[2] * q, v0[3] * q), umma::pk_bf16(v1[0] * q, v1[1] * q), umma::pk_bf16(v1[2] * q, v1[3] * q),
                    umma::pk_bf16(t, mf[0]));
  u[2] = make_uint4(umma::pk_bf16(mf[1], mf[2]), umma::pk_bf16(mf[3], mf[4]), umma::pk_bf16(mf[5], mf[6]), umma::pk_bf16(mf[7], mf[8]));
  u[3] = make_uint4(umma::pk_bf16(fs.x * qf, fs.y * qf), umma::pk_bf16(fs.z * qf, fs.w * qf), 0u, 0u);
  const int Hs = Hp >> 1, Ws = Wp >> 1;
  size_t o;
  if (phase) o = ((((size_t)n * 4 + (Y & 1) * 2 + (X & 1)) * (Hs >> 1) + (Y >> 1)) * (Ws >> 1) + (X >> 1)) * 32;
  else o = (((size_t)n * Hs + Y) * Ws + X) * 32;
  st_v8(out + o, u[0], u[1]);
  st_v8(out + o + 16, u[2], u[3]);
}
static int launch_gather_s2(vfi_handle* h, const uint8_t* imgs, const bf16* feats, const LowSet& lows, const float* tsteps_dev,
                            bf16* out, int B, int I1, int T, int Hp, int Wp, bool phase, cudaStream_t s) {
  GcRows grows;
  const size_t smem = gc_stage_bytes(lows, &grows, TL_W);
  const int Tr = h->opt_t_fast ? T : 1;
  const dim3 grid(Wp / TL_W * Tr, Hp / TL_H, B * T / Tr);
  gather_s2_kernel<<<grid, G2_THREADS, smem, s>>>(imgs, feats, lows, grows, tsteps_dev, out, B, I1, T, Hp, Wp, phase ? 1 : 0, Tr);
  VFI_LAUNCHED(h);
  return 0;
}

// ------------------------------------------------------------------------------------------
// Block-input gather at s >= 4 (bf16 mode, LowPlanes levels): gather_kernel<bf16, false> with the per-call tap tables
// instead of per-thread source-index arithmetic, 32-bit indexing, and mask / feat fetched once per output with the
// mean weights of the 2 x 2 sampled positions (which share their source cell: the last level is at least 2s coarse).
// ------------------------------------------------------------------------------------------
// Block 3 at scale 1.0 (s = 2): block-input gather FUSED with conv0's first stride-2 convolution (28 -> 32 channels), the
// way gather_conv0a_p_kernel does it for the finest block.  The 28-channel half-resolution block input (64 B per pixel,
// 3 GB written and read back per 90 items of 1080p) never goes to HBM.
//   tile     : 8 x 16 conv outputs (one M = 128 MMA tile) <- 17 x 33 half-resolution input pixels (stride 2, pad 1) <- the
//              34 x 66 full-resolution pixels whose four warped samples each of them averages (gather_s2_kernel's arithmetic)
//   A (smem) : conv_s2.cu's four phase-plane windows [Cin/8][plane][17 rows][9 cols][8 ch], written straight by the gather
//   B (smem) : [tap][Cin/8][32][8] bf16 (18 KB), resident; D: 32 TMEM columns
//   persistent, 288 threads (561 pixels in two rounds), 2 CTAs per SM; warps 0-3 issue the MMAs and run the epilogue of tile
//   i while warps 4-8 stage the flow rows of tile i+1.
#define G3_THREADS 288
#ifndef G3_UNROLL
#define G3_UNROLL 2
#endif
constexpr int kG3Unroll = G3_UNROLL;   // samples of a pixel in flight together (the kernel runs 18 warps per SM: ILP instead of occupancy)
#define G3_RW 17          // region width  (half-resolution pixels)
#define G3_RH 33          // region height
#define G3_FW 34          // flow region width  (full-resolution pixels)
#define G3_FH 66          // flow region height
#define G3_PW 9           // phase-plane window width
#define G3_PLANE 156      // plane stride (16-byte units): >= 17 * 9, == 4 mod 8
#define G3_CS 626         // chunk stride: >= 4 * G3_PLANE, == 2 mod 8
#define G3_A_BYTES (4 * G3_CS * 16)
#define G3_W_BYTES (9 * 32 * 32 * 2)
#define G3_TAB (5 * G3_FH)
__host__ __device__ __forceinline__ int g3_rows(int s) { return G3_FH / s + 3; }   // source rows a 66-row span can touch (+1 spare)
static size_t g3_stage_bytes(const LowSet& lows, GcRows* R) {
  int rows = 0;
  for (int k = 0; k < 5; k++) {
    R->start[k] = rows;
    if (k < lows.n) rows += g3_rows(lows.s[k]);
  }
  R->start[5] = rows;
  return ((size_t)G3_TAB + (size_t)rows * G3_FW) * sizeof(float4);
}
// flat_stage for a G3_FW x G3_FH region with origin (ox, oy): per-level row table (one entry per region row) followed by the
// x-interpolated, pre-scaled flow rows of every coarse level (G3_FW float4 per row)
template <int NT>
__device__ __forceinline__ void rect_stage(const LowSet& L, const GcRows& R, int n, int ox, int oy, int Hp, int Wp, float4* sm, int tid) {
  RowTap* ytab = reinterpret_cast<RowTap*>(sm);
  const float oyf = (float)max(oy, 0) + 0.5f;
  for (int e = tid; e < L.n * G3_FH; e += NT) {
    const int k = e / G3_FH, yl = e - k * G3_FH;
    const Tap ty = ld_tap(L.yt[k] + min(max(oy + yl, 0), Hp - 1));
    const int r_lo = (int)fmaxf(L.rs[k] * oyf - 0.5f, 0.f);
    const int off = G3_TAB + R.start[k] * G3_FW;
    RowTap t;
    t.o0 = (off + (ty.i0 - r_lo) * G3_FW) * 16; t.o1 = (off + (ty.i1 - r_lo) * G3_FW) * 16;  // byte offsets
    t.hy = ty.h; t.ly = ty.l;
    ytab[e] = t;
  }
  const int total = R.start[5] * G3_FW;
#pragma unroll 2
  for (int e = tid; e < total; e += NT) {
    const int u = e / G3_FW, col = e - u * G3_FW;
    int k = 0;
#pragma unroll
    for (int j = 1; j < 5; j++) k += (j < L.n && u >= R.start[j]) ? 1 : 0;
    const int Hl = L.Hl[k], Wl = L.Wl[k], c4 = L.c4[k];
    const int r_lo = (int)fmaxf(L.rs[k] * oyf - 0.5f, 0.f);
    const int ry = min(r_lo + u - R.start[k], Hl - 1);
    const Tap tx = ld_tap(L.xt[k] + min(max(ox + col, 0), Wp - 1));
    const float4* row = reinterpret_cast<const float4*>(L.p[k]) + ((size_t)n * Hl + ry) * Wl * c4;
    const float4 a = __ldg(row + tx.i0 * c4), b = __ldg(row + tx.i1 * c4);
    const float fs = L.fs[k];
    sm[G3_TAB + e] = make_float4((tx.h * a.x + tx.l * b.x) * fs, (tx.h * a.y + tx.l * b.y) * fs,
                                 (tx.h * a.z + tx.l * b.z) * fs, (tx.h * a.w + tx.l * b.w) * fs);
  }
}
__device__ __forceinline__ float4 rect_eval(const LowSet& L, const float4* sm, int lx, int ly) {
  const RowTap* ytab = reinterpret_cast<const RowTap*>(sm);
  float4 fl = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
  for (int k = 0; k < 5; k++) {
    if (k < L.n) {
      const RowTap t = ytab[k * G3_FH + ly];
      const char* col = reinterpret_cast<const char*>(sm + lx);
      const float4 a = *reinterpret_cast<const float4*>(col + t.o0), c = *reinterpret_cast<const float4*>(col + t.o1);
      fl.x = fmaf(t.ly, c.x, fmaf(t.hy, a.x, fl.x)); fl.y = fmaf(t.ly, c.y, fmaf(t.hy, a.y, fl.y));
      fl.z = fmaf(t.ly, c.z, fmaf(t.hy, a.z, fl.z)); fl.w = fmaf(t.ly, c.w, fmaf(t.hy, a.w, fl.w));
    }
  }
  return fl;
}

__device__ __forceinline__ uint32_t g3_pix_index(int ly, int lx) {
  // region-local (ly, lx) in [0, 33) x [0, 17): half-resolution pixel (hy0 - 1 + ly, hx0 - 1 + lx); hy0, hx0 are even
  const int py = (ly + 1) & 1, px = (lx + 1) & 1;
  const int pr = (ly + 1) >> 1, pc = (lx + 1) >> 1;
  return (uint32_t)((py * 2 + px) * G3_PLANE + pr * G3_PW + pc);
}
__global__ void __launch_bounds__(G3_THREADS, 2)
gather_s2_conv0a_p_kernel(const uint8_t* __restrict__ imgs, const bf16* __restrict__ feats, const __grid_constant__ LowSet lows,
                          const float* __restrict__ tsteps, const bf16* __restrict__ wpack, const float* __restrict__ bias,
                          bf16* __restrict__ out, const __grid_constant__ GcRows grows, int B, int I1, int Tn, int Hp, int Wp,
                          float slope, int stage_bytes, int tiles_x, int tiles_y, int num_tiles) {
  using namespace umma;
  extern __shared__ float4 tl_sm[];  // row table + staged flow rows | A tile | weights
  __shared__ uint64_t mma_bar;
  __shared__ uint32_t tmem_slot;
  uint8_t* a_s = reinterpret_cast<uint8_t*>(tl_sm) + stage_bytes;
  uint8_t* w_s = a_s + G3_A_BYTES;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) { bar_init(&mma_bar, 1); fence_barrier_init(); }
  if (warp == 0) tmem_alloc(&tmem_slot, 32);
  for (int i = threadIdx.x; i < G3_W_BYTES / 16; i += G3_THREADS)
    reinterpret_cast<uint4*>(w_s)[i] = __ldg(reinterpret_cast<const uint4*>(wpack) + i);
  int tile = blockIdx.x;
  int ctx = tile % tiles_x, cty = (tile / tiles_x) % tiles_y, cn = tile / (tiles_x * tiles_y);
  const int g = (int)gridDim.x, gsx = g % tiles_x, gsy = (g / tiles_x) % tiles_y, gsn = g / (tiles_x * tiles_y);
  // full-resolution origin of tile (tx, ty)'s flow region: two pixels left of / above its 32 x 64 full-resolution footprint
  if (tile < num_tiles) rect_stage<G3_THREADS>(lows, grows, cn, ctx * 32 - 2, cty * 64 - 2, Hp, Wp, tl_sm, (int)threadIdx.x);
  fence_async_smem();
  tc_before();
  __syncthreads();
  tc_after();
  const uint32_t tmem_base = tmem_slot;
  const int plane = Hp * Wp;
  const uint32_t* img32 = reinterpret_cast<const uint32_t*>(imgs);
  const int kl = lows.n - 1, mWl = lows.Wl[kl];
  const bf16 *mfp = lows.mf[kl], *f7p = lows.f7[kl];
  const Tap *mxt = lows.xt[kl], *myt = lows.yt[kl];
  for (int it = 0; tile < num_tiles; tile += g, it++) {
    const int hx0 = ctx * 16, hy0 = cty * 32, n = cn;      // half-resolution tile origin
    ctx += gsx; cty += gsy; cn += gsn;                       // the next tile of this CTA
    if (ctx >= tiles_x) { ctx -= tiles_x; cty++; }
    if (cty >= tiles_y) { cty -= tiles_y; cn++; }
    const int b = n / Tn;
    const float t = tsteps[n - b * Tn];
    const int base0 = b * plane, base1 = (I1 + b) * plane;
    const int mtb = n * lows.Hl[kl] * mWl;
#pragma unroll 1
    for (int e = threadIdx.x; e < G3_RW * G3_RH; e += G3_THREADS) {
      const int ly = e / G3_RW, lx = e - ly * G3_RW;
      const int hx = hx0 - 1 + lx, hy = hy0 - 1 + ly;        // half-resolution pixel
      uint4 u[4];
#pragma unroll
      for (int j = 0; j < 4; j++) u[j] = make_uint4(0u, 0u, 0u, 0u);
      if (hx >= 0 && hy >= 0) {                               // else: the convolution's zero padding
        const int gx = 2 * hx, gy = 2 * hy;                   // top-left of the pixel's 2 x 2 full-resolution positions
        float mf[9];
        {
          const Tap xa = ld_tap(mxt + gx), xb = ld_tap(mxt + gx + 1), ya = ld_tap(myt + gy), yb = ld_tap(myt + gy + 1);
          const float hxw = 0.5f * (xa.h + xb.h), lxw = 0.5f * (xa.l + xb.l), hyw = 0.5f * (ya.h + yb.h), lyw = 0.5f * (ya.l + yb.l);
          const int r0 = mtb + ya.i0 * mWl, r1 = mtb + ya.i1 * mWl;
          const float w00 = hyw * hxw, w01 = hyw * lxw, w10 = lyw * hxw, w11 = lyw * lxw;
          float a[9], c[9];
          mf9_load(mfp, f7p, r0 + xa.i0, a);
          mf9_load(mfp, f7p, r0 + xa.i1, c);
#pragma unroll
          for (int i = 0; i < 9; i++) mf[i] = w00 * a[i] + w01 * c[i];
          mf9_load(mfp, f7p, r1 + xa.i0, a);
          mf9_load(mfp, f7p, r1 + xa.i1, c);
#pragma unroll
          for (int i = 0; i < 9; i++) mf[i] += w10 * a[i] + w11 * c[i];
        }
        float c0[3] = {0.f, 0.f, 0.f}, c1[3] = {0.f, 0.f, 0.f}, v0[4] = {0.f, 0.f, 0.f, 0.f}, v1[4] = {0.f, 0.f, 0.f, 0.f};
        float4 fs = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll kG3Unroll
        for (int d = 0; d < 4; d++) {                         // same visiting order and sums as gather_s2_kernel
          const int dx = d & 1, dy = d >> 1;
          const float4 fl = rect_eval(lows, tl_sm, 2 * lx + dx, 2 * ly + dy);
          const Bil b0 = bil_setup((float)(gx + dx) + fl.x, (float)(gy + dy) + fl.y, Wp, Hp, base0);
          const Bil b1 = bil_setup((float)(gx + dx) + fl.z, (float)(gy + dy) + fl.w, Wp, Hp, base1);
          float r[4];
          sample_rgbx(img32, b0, r);
          c0[0] += r[0]; c0[1] += r[1]; c0[2] += r[2];
          sample_rgbx(img32, b1, r);
          c1[0] += r[0]; c1[1] += r[1]; c1[2] += r[2];
          sample_feat_px<bf16>(feats, b0, r);
          v0[0] += r[0]; v0[1] += r[1]; v0[2] += r[2]; v0[3] += r[3];
          sample_feat_px<bf16>(feats, b1, r);
          v1[0] += r[0]; v1[1] += r[1]; v1[2] += r[2]; v1[3] += r[3];
          fs.x += fl.x; fs.y += fl.y; fs.z += fl.z; fs.w += fl.w;
        }
        const float q = 0.25f, qf = 0.125f;  // mean of the four positions; the flow also scales by 1/s = 1/2
        u[0] = make_uint4(pk_bf16(c0[0] * q, c0[1] * q), pk_bf16(c0[2] * q, c1[0] * q), pk_bf16(c1[1] * q, c1[2] * q), pk_bf16(v0[0] * q, v0[1] * q));
        u[1] = make_uint4(pk_bf16(v0[2] * q, v0[3] * q), pk_bf16(v1[0] * q, v1[1] * q), pk_bf16(v1[2] * q, v1[3] * q), pk_bf16(t, mf[0]));
        u[2] = make_uint4(pk_bf16(mf[1], mf[2]), pk_bf16(mf[3], mf[4]), pk_bf16(mf[5], mf[6]), pk_bf16(mf[7], mf[8]));
        u[3] = make_uint4(pk_bf16(fs.x * qf, fs.y * qf), pk_bf16(fs.z * qf, fs.w * qf), 0u, 0u);
      }
      const uint32_t pix = g3_pix_index(ly, lx);
#pragma unroll
      for (int j = 0; j < 4; j++) *reinterpret_cast<uint4*>(a_s + ((size_t)j * G3_CS + pix) * 16) = u[j];
    }
    fence_async_smem();
    __syncthreads();               // the A tile is complete; the staging area is free again
    if (warp == 0) {
      const uint32_t leader = elect_one();
      constexpr uint32_t IDESC = idesc_bf16_m128(32);
      const uint64_t ad0 = desc_ns(s_u32(a_s), G3_CS * 16, G3_PW * 16);
      const uint64_t bd0 = desc_ns(s_u32(w_s), 32 * 16, 128);
#pragma unroll
      for (int tap = 0; tap < 9; tap++) {
#pragma unroll
        for (int k = 0; k < 2; k++) {
          const int ky = tap / 3, kx = tap % 3;
          const int pl = (ky != 1 ? 2 : 0) + (kx != 1 ? 1 : 0);
          const uint32_t a_off = (uint32_t)(pl * G3_PLANE + (ky != 0 ? 1 : 0) * G3_PW + (kx != 0 ? 1 : 0) + 2 * k * G3_CS);
          const uint32_t b_off = (uint32_t)((tap * 4 + 2 * k) * 32);
          mma_bf16_elect(leader, tmem_base, ad0 + a_off, bd0 + b_off, IDESC, (tap | k) ? 1u : 0u);
        }
      }
      if (leader) commit_to(&mma_bar);
      __syncwarp();
    }
    if (warp < 4) {
      bar_wait(&mma_bar, (uint32_t)(it & 1));
      tc_after();
      const int m = warp * 32 + lane;
      const int Ho = Hp >> 2, Wo = Wp >> 2;                   // conv0[0]'s output grid (quarter resolution)
      const int oy = (hy0 >> 1) + (m >> 3), ox = (hx0 >> 1) + (m & 7);
      float v[32];
      ld16(tmem_base + ((uint32_t)(warp * 32) << 16), v);
      ld16(tmem_base + ((uint32_t)(warp * 32) << 16) + 16, v + 16);
#pragma unroll
      for (int c = 0; c < 32; c++) {
        const float x = v[c] + __ldg(bias + c);
        v[c] = x > 0.f ? x : x * slope;
      }
      // conv0's second conv is stride 2 too: write its four phase planes [N, 4, Ho/2, Wo/2, 32]
      const size_t opix = ((((size_t)n * 4 + (oy & 1) * 2 + (ox & 1)) * (Ho >> 1) + (oy >> 1)) * (Wo >> 1) + (ox >> 1));
      uint4 o[4];
#pragma unroll
      for (int j = 0; j < 4; j++)
        o[j] = make_uint4(pk_bf16(v[8 * j], v[8 * j + 1]), pk_bf16(v[8 * j + 2], v[8 * j + 3]), pk_bf16(v[8 * j + 4], v[8 * j + 5]),
                          pk_bf16(v[8 * j + 6], v[8 * j + 7]));
      st_v8(out + opix * 32, o[0], o[1]);
      st_v8(out + opix * 32 + 16, o[2], o[3]);
    }
    if (warp >= 4 && tile + g < num_tiles)
      rect_stage<G3_THREADS - 128>(lows, grows, cn, ctx * 32 - 2, cty * 64 - 2, Hp, Wp, tl_sm, (int)threadIdx.x - 128);
    tc_before();
    __syncthreads();               // next tile staged; this tile's accumulator read out
    tc_after();
  }
  if (warp == 0) tmem_dealloc(tmem_base, 32);
}
// Lb0 = block 3's conv0[0] (28 -> 32, stride 2).  Output: its activation as the phase planes of [N, Hp/4, Wp/4, 32].
int k_gather_s2_conv0a(vfi_handle* h, const Layer& Lb0, const uint8_t* imgs, const void* feats, const LowSet& lows,
                       const float* tsteps_dev, bf16* out, int B, int I1, int T, int Hp, int Wp, cudaStream_t s) {
  ProfScope _ps(h, PC_GATHER, s);
  if (!Lb0.ws2 || Lb0.cin_pad != 32 || Lb0.cout != 32) VFI_FAIL(h, VFI_E_STATE, "gather_s2_conv0a: layer not packed");
  if (Hp % 64 || Wp % 32) VFI_FAIL(h, VFI_E_ARG, "gather_s2_conv0a: padded size must be a multiple of 64 x 32");
  for (int k = 0; k < lows.n; k++)
    if (lows.s[k] < 4) VFI_FAIL(h, VFI_E_ARG, "gather_s2_conv0a: coarse levels only");
  if (lows.n < 1 || !lows.mf[lows.n - 1] || !lows.xt[0]) VFI_FAIL(h, VFI_E_ARG, "gather_s2_conv0a: needs tap tables and plane-stored coarse levels");
  GcRows grows;
  const size_t un = (g3_stage_bytes(lows, &grows) + 127) & ~(size_t)127;
  const size_t smem = un + G3_A_BYTES + G3_W_BYTES;
  if (smem > 110 * 1024) VFI_FAIL(h, VFI_E_ARG, "gather_s2_conv0a: shared memory budget");
  VFI_CUDA(h, cudaFuncSetAttribute(gather_s2_conv0a_p_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 110 * 1024));
  const int tiles_x = Wp / 32, tiles_y = Hp / 64, num_tiles = tiles_x * tiles_y * B * T;
  gather_s2_conv0a_p_kernel<<<std::min(num_tiles, 2 * h->sm_count), G3_THREADS, smem, s>>>(
      imgs, (const bf16*)feats, lows, tsteps_dev, Lb0.ws2, Lb0.bias, out, grows, B, I1, T, Hp, Wp, h->slope, (int)un, tiles_x, tiles_y, num_tiles);
  VFI_LAUNCHED(h);
  return 0;
}

__global__ void __launch_bounds__(128) gather_lo_kernel(const uint8_t* __restrict__ imgs, const bf16* __restrict__ feats,
                                                        const __grid_constant__ LowSet lows, const float* __restrict__ tsteps,
                                                        bf16* __restrict__ out, int B, int I1, int Tn, int Hp, int Wp, int s,
                                                        int phase, int Tr) {
  const int Hs = Hp / s, Ws = Wp / s;
  const size_t i0 = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i0 >= (size_t)B * Tn * Hs * Ws) return;
  // thread order: x, then (Tr = T) the timestep, then y, then the pair -- a row of all T items before the next row, so the items
  // of a pair share their image reads in L2; Tr = 1 is the item-major order.  i = the item-major index the output uses
  const int X = (int)(i0 % Ws);
  const size_t r = i0 / Ws, r2 = r / Tr;
  const int Y = (int)(r2 % Hs), n = (int)(r2 / Hs) * Tr + (int)(r - r2 * Tr);
  const size_t i = ((size_t)n * Hs + Y) * Ws + X;
  const int b = n / Tn;
  const float t = tsteps[n - b * Tn];
  const int plane = Hp * Wp;
  const uint32_t* img32 = reinterpret_cast<const uint32_t*>(imgs);
  const int base0 = b * plane, base1 = (I1 + b) * plane;
  const int bx = s * X + s / 2 - 1, by = s * Y + s / 2 - 1;
  float mf[9];
  {
    const int kl = lows.n - 1, Wl = lows.Wl[kl];
    const Tap xa = ld_tap(lows.xt[kl] + bx), xb = ld_tap(lows.xt[kl] + bx + 1);
    const Tap ya = ld_tap(lows.yt[kl] + by), yb = ld_tap(lows.yt[kl] + by + 1);
    const float hx = 0.5f * (xa.h + xb.h), lx = 0.5f * (xa.l + xb.l), hy = 0.5f * (ya.h + yb.h), ly = 0.5f * (ya.l + yb.l);
    const int r0 = (n * lows.Hl[kl] + ya.i0) * Wl, r1 = (n * lows.Hl[kl] + ya.i1) * Wl;
    const float w00 = hy * hx, w01 = hy * lx, w10 = ly * hx, w11 = ly * lx;
    float a[9], c[9];
    mf9_load(lows.mf[kl], lows.f7[kl], r0 + xa.i0, a);
    mf9_load(lows.mf[kl], lows.f7[kl], r0 + xa.i1, c);
#pragma unroll
    for (int j = 0; j < 9; j++) mf[j] = w00 * a[j] + w01 * c[j];
    mf9_load(lows.mf[kl], lows.f7[kl], r1 + xa.i0, a);
    mf9_load(lows.mf[kl], lows.f7[kl], r1 + xa.i1, c);
#pragma unroll
    for (int j = 0; j < 9; j++) mf[j] += w10 * a[j] + w11 * c[j];
  }
  float c0[3] = {0.f, 0.f, 0.f}, c1[3] = {0.f, 0.f, 0.f}, v0[4] = {0.f, 0.f, 0.f, 0.f}, v1[4] = {0.f, 0.f, 0.f, 0.f};
  float4 fs = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 1
  for (int d = 0; d < 4; d++) {
    const int x = bx + (d & 1), y = by + (d >> 1);
    float4 fl = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int k = 0; k < 5; k++) {
      if (k < lows.n) {
        const Tap tx = ld_tap(lows.xt[k] + x), ty = ld_tap(lows.yt[k] + y);
        const float4* fp = reinterpret_cast<const float4*>(lows.p[k]);
        const int r0 = (n * lows.Hl[k] + ty.i0) * lows.Wl[k], r1 = (n * lows.Hl[k] + ty.i1) * lows.Wl[k];
        const float4 a = __ldg(fp + r0 + tx.i0), bq = __ldg(fp + r0 + tx.i1), c = __ldg(fp + r1 + tx.i0), dq = __ldg(fp + r1 + tx.i1);
        const float f = lows.fs[k];
        fl.x += (ty.h * (tx.h * a.x + tx.l * bq.x) + ty.l * (tx.h * c.x + tx.l * dq.x)) * f;
        fl.y += (ty.h * (tx.h * a.y + tx.l * bq.y) + ty.l * (tx.h * c.y + tx.l * dq.y)) * f;
        fl.z += (ty.h * (tx.h * a.z + tx.l * bq.z) + ty.l * (tx.h * c.z + tx.l * dq.z)) * f;
        fl.w += (ty.h * (tx.h * a.w + tx.l * bq.w) + ty.l * (tx.h * c.w + tx.l * dq.w)) * f;
      }
    }
    const Bil b0 = bil_setup((float)x + fl.x, (float)y + fl.y, Wp, Hp, base0);
    const Bil b1 = bil_setup((float)x + fl.z, (float)y + fl.w, Wp, Hp, base1);
    float r[4];
    sample_rgbx(img32, b0, r);
    c0[0] += r[0]; c0[1] += r[1]; c0[2] += r[2];
    sample_rgbx(img32, b1, r);
    c1[0] += r[0]; c1[1] += r[1]; c1[2] += r[2];
    sample_feat_px<bf16>(feats, b0, r);
    v0[0] += r[0]; v0[1] += r[1]; v0[2] += r[2]; v0[3] += r[3];
    sample_feat_px<bf16>(feats, b1, r);
    v1[0] += r[0]; v1[1] += r[1]; v1[2] += r[2]; v1[3] += r[3];
    fs.x += fl.x; fs.y += fl.y; fs.z += fl.z; fs.w += fl.w;
  }
  const float q = 0.25f, qf = 0.25f / (float)s;  // mean of the four positions; the flow also scales by 1/s
  uint4 u[4];
  u[0] = make_uint4(umma::pk_bf16(c0[0] * q, c0[1] * q), umma::pk_bf16(c0[2] * q, c1[0] * q), umma::pk_bf16(c1[1] * q, c1[2] * q),
                    umma::pk_bf16(v0[0] * q, v0[1] * q));
  u[1] = make_uint4(umma::pk_bf16(v0[2] * q, v0[3] * q), umma::pk_bf16(v1[0] * q, v1[1] * q), umma::pk_bf16(v1[2] * q, v1[3] * q),
                    umma::pk_bf16(t, mf[0]));
  u[2] = make_uint4(umma::pk_bf16(mf[1], mf[2]), umma::pk_bf16(mf[3], mf[4]), umma::pk_bf16(mf[5], mf[6]), umma::pk_bf16(mf[7], mf[8]));
  u[3] = make_uint4(umma::pk_bf16(fs.x * qf, fs.y * qf), umma::pk_bf16(fs.z * qf, fs.w * qf), 0u, 0u);
  size_t o;
  if (phase) o = ((((size_t)n * 4 + (Y & 1) * 2 + (X & 1)) * (Hs >> 1) + (Y >> 1)) * (Ws >> 1) + (X >> 1)) * 32;
  else o = i * 32;
  st_v8(out + o, u[0], u[1]);
  st_v8(out + o + 16, u[2], u[3]);
}
static int launch_gather_lo(vfi_handle* h, const uint8_t* imgs, const bf16* feats, const LowSet& lows, const float* tsteps_dev,
                            bf16* out, int B, int I1, int T, int Hp, int Wp, int s_int, bool phase, cudaStream_t s) {
  const size_t total = (size_t)B * T * (Hp / s_int) * (Wp / s_int);
  gather_lo_kernel<<<cdiv(total, 128), 128, 0, s>>>(imgs, feats, lows, tsteps_dev, out, B, I1, T, Hp, Wp, s_int, phase ? 1 : 0,
                                                    h->opt_t_fast ? T : 1);
  VFI_LAUNCHED(h);
  return 0;
}
