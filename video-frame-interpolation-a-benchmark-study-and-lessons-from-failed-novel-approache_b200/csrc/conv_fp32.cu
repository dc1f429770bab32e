// conv_fp32.cu -- CUDA-core direct convolutions with fp32 accumulation.
// They are (1) the whole conv path of VFI_PRECISION_FP32 (the 1e-4 parity mode of north_star),
// (2) the Head's 16->4 ConvTranspose in both precisions (N = 4 is no tensor-core shape), and
// (3) the on-GPU cross-check of the tcgen05 kernels in tests (vfi_debug_conv).
// Semantics: Conv2d(k3, s1|s2, p1) [+ *beta + residual] [+ LeakyReLU]; ConvTranspose2d(k4, s2, p1)
// [+ PixelShuffle(2)]  (SURVEY.md Appendix A.2).
#include "common.cuh"

template <typename T> struct Elt;
template <> struct Elt<float> {
  static __device__ __forceinline__ void load4(const float* p, float* v) {
    float4 a = __ldg(reinterpret_cast<const float4*>(p));
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
  }
  static __device__ __forceinline__ void store4(float* p, const float* v) {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  }
};
template <> struct Elt<bf16> {
  static __device__ __forceinline__ void load4(const bf16* p, float* v) {
    uint2 a = __ldg(reinterpret_cast<const uint2*>(p));
    float2 f0 = __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&a.x));
    float2 f1 = __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&a.y));
    v[0] = f0.x; v[1] = f0.y; v[2] = f1.x; v[3] = f1.y;
  }
  static __device__ __forceinline__ void store4(bf16* p, const float* v) {
    __nv_bfloat162 a = __floats2bfloat162_rn(v[0], v[1]), b = __floats2bfloat162_rn(v[2], v[3]);
    uint2 u;
    u.x = *reinterpret_cast<uint32_t*>(&a); u.y = *reinterpret_cast<uint32_t*>(&b);
    *reinterpret_cast<uint2*>(p) = u;
  }
};

template <typename T>
__global__ void conv3_direct_kernel(const T* __restrict__ in, const float* __restrict__ w,
                                    const float* __restrict__ bias, const float* __restrict__ beta,
                                    const T* __restrict__ res, T* __restrict__ out, int N, int Hin, int Win, int Ho,
                                    int Wo, int cin_pad, int cout_pad, int stride, int act, float slope) {
  const int G = cout_pad >> 2;
  const int ppb = blockDim.x / G;
  const int g = threadIdx.x % G, pl = threadIdx.x / G;
  if (pl >= ppb) return;
  size_t pix = (size_t)blockIdx.x * ppb + pl;
  if (pix >= (size_t)N * Ho * Wo) return;
  int x = (int)(pix % Wo), y = (int)((pix / Wo) % Ho), n = (int)(pix / ((size_t)Wo * Ho));
  float acc[4];
  {
    float4 b4 = __ldg(reinterpret_cast<const float4*>(bias) + g);
    acc[0] = b4.x; acc[1] = b4.y; acc[2] = b4.z; acc[3] = b4.w;
  }
  for (int r = 0; r < 3; r++) {
    int iy = y * stride + r - 1;
    if (iy < 0 || iy >= Hin) continue;
    for (int q = 0; q < 3; q++) {
      int ix = x * stride + q - 1;
      if (ix < 0 || ix >= Win) continue;
      const T* ip = in + (((size_t)n * Hin + iy) * Win + ix) * cin_pad;
      const float4* wp = reinterpret_cast<const float4*>(w + (size_t)(r * 3 + q) * cin_pad * cout_pad) + g;
      for (int ci = 0; ci < cin_pad; ci += 4) {
        float v[4];
        Elt<T>::load4(ip + ci, v);
#pragma unroll
        for (int k = 0; k < 4; k++) {
          float4 w4 = __ldg(wp + (size_t)(ci + k) * G);
          acc[0] = fmaf(v[k], w4.x, acc[0]); acc[1] = fmaf(v[k], w4.y, acc[1]);
          acc[2] = fmaf(v[k], w4.z, acc[2]); acc[3] = fmaf(v[k], w4.w, acc[3]);
        }
      }
    }
  }
  if (beta) {
    float4 be = __ldg(reinterpret_cast<const float4*>(beta) + g);
    float rv[4];
    Elt<T>::load4(res + pix * cout_pad + 4 * g, rv);
    acc[0] = acc[0] * be.x + rv[0]; acc[1] = acc[1] * be.y + rv[1];
    acc[2] = acc[2] * be.z + rv[2]; acc[3] = acc[3] * be.w + rv[3];
  }
  if (act) {
#pragma unroll
    for (int k = 0; k < 4; k++) acc[k] = acc[k] > 0.f ? acc[k] : acc[k] * slope;
  }
  Elt<T>::store4(out + pix * cout_pad + 4 * g, acc);
}

int k_conv3_direct(vfi_handle* h, const Layer& L, const void* in, const void* residual, void* out, int N, int Hin,
                   int Win, bool bf, float slope, cudaStream_t s) {
  ProfScope _ps(h, PC_CONV_DIRECT, s);
  int Ho = L.stride == 2 ? Hin / 2 : Hin, Wo = L.stride == 2 ? Win / 2 : Win;
  int G = L.cout_pad / 4;
  int threads = 256;
  int ppb = threads / G;
  if (ppb < 1) VFI_FAIL(h, VFI_E_ARG, "conv3_direct: cout_pad %d too large", L.cout_pad);
  size_t pixels = (size_t)N * Ho * Wo;
  int grid = cdiv(pixels, ppb);
  const float* beta = L.has_beta ? L.beta : nullptr;
  if (beta && !residual) VFI_FAIL(h, VFI_E_ARG, "conv3_direct: ResConv layer needs a residual");
  if (bf)
    conv3_direct_kernel<bf16><<<grid, threads, 0, s>>>((const bf16*)in, L.w32, L.bias, beta, (const bf16*)residual,
                                                      (bf16*)out, N, Hin, Win, Ho, Wo, L.cin_pad, L.cout_pad,
                                                      L.stride, L.act ? 1 : 0, slope);
  else
    conv3_direct_kernel<float><<<grid, threads, 0, s>>>((const float*)in, L.w32, L.bias, beta,
                                                       (const float*)residual, (float*)out, N, Hin, Win, Ho, Wo,
                                                       L.cin_pad, L.cout_pad, L.stride, L.act ? 1 : 0, slope);
  VFI_LAUNCHED(h);
  return 0;
}

template <typename T>
__global__ void deconv4_direct_kernel(const T* __restrict__ in, const float* __restrict__ w,
                                      const float* __restrict__ bias, void* __restrict__ out, int N, int Hin,
                                      int Win, int cin_pad, int cout_pad, int mode) {
  const int G = cout_pad >> 2;
  const int ppb = blockDim.x / G;
  const int g = threadIdx.x % G, pl = threadIdx.x / G;
  if (pl >= ppb) return;
  const int Ho = 2 * Hin, Wo = 2 * Win;
  size_t pix = (size_t)blockIdx.x * ppb + pl;
  if (pix >= (size_t)N * Ho * Wo) return;
  int ox = (int)(pix % Wo), oy = (int)((pix / Wo) % Ho), n = (int)(pix / ((size_t)Wo * Ho));
  float acc[4];
  {
    float4 b4 = __ldg(reinterpret_cast<const float4*>(bias) + g);
    acc[0] = b4.x; acc[1] = b4.y; acc[2] = b4.z; acc[3] = b4.w;
  }
  for (int ky = 0; ky < 4; ky++) {
    int ty = oy + 1 - ky;
    if (ty < 0 || (ty & 1)) continue;
    int iy = ty >> 1;
    if (iy >= Hin) continue;
    for (int kx = 0; kx < 4; kx++) {
      int tx = ox + 1 - kx;
      if (tx < 0 || (tx & 1)) continue;
      int ix = tx >> 1;
      if (ix >= Win) continue;
      const T* ip = in + (((size_t)n * Hin + iy) * Win + ix) * cin_pad;
      const float4* wp = reinterpret_cast<const float4*>(w + (size_t)(ky * 4 + kx) * cin_pad * cout_pad) + g;
      for (int ci = 0; ci < cin_pad; ci += 4) {
        float v[4];
        Elt<T>::load4(ip + ci, v);
#pragma unroll
        for (int k = 0; k < 4; k++) {
          float4 w4 = __ldg(wp + (size_t)(ci + k) * G);
          acc[0] = fmaf(v[k], w4.x, acc[0]); acc[1] = fmaf(v[k], w4.y, acc[1]);
          acc[2] = fmaf(v[k], w4.z, acc[2]); acc[3] = fmaf(v[k], w4.w, acc[3]);
        }
      }
    }
  }
  if (mode == 0) {
    Elt<T>::store4(reinterpret_cast<T*>(out) + pix * cout_pad + 4 * g, acc);
  } else {
    // PixelShuffle(2): channel j = 4*cc + 2*i + jj of pixel (oy,ox) -> channel cc of pixel (2oy+i, 2ox+jj)
    float* o = reinterpret_cast<float*>(out);
    const int H4 = 2 * Ho, W4 = 2 * Wo;
#pragma unroll
    for (int k = 0; k < 4; k++) {
      int i = k >> 1, jj = k & 1;
      o[(((size_t)n * H4 + 2 * oy + i) * W4 + 2 * ox + jj) * 16 + g] = acc[k];
    }
  }
}

int k_deconv4_direct(vfi_handle* h, const Layer& L, const void* in, void* out, int N, int Hin, int Win, bool bf,
                     int mode, cudaStream_t s) {
  ProfScope _ps(h, PC_DECONV_DIRECT, s);
  int G = L.cout_pad / 4;
  int threads = 256;
  int ppb = threads / G;
  size_t pixels = (size_t)N * 4 * Hin * Win;
  int grid = cdiv(pixels, ppb);
  if (bf)
    deconv4_direct_kernel<bf16><<<grid, threads, 0, s>>>((const bf16*)in, L.w32, L.bias, out, N, Hin, Win, L.cin_pad,
                                                        L.cout_pad, mode);
  else
    deconv4_direct_kernel<float><<<grid, threads, 0, s>>>((const float*)in, L.w32, L.bias, out, N, Hin, Win,
                                                         L.cin_pad, L.cout_pad, mode);
  VFI_LAUNCHED(h);
  return 0;
}
