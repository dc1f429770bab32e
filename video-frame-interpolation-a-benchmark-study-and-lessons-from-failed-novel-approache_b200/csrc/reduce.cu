// reduce.cu -- warp-shuffle reduction kernels (north_star kernel group (c)): PSNR, SSIM (full-res
// metric and quarter-res scene-cut score), flow-magnitude statistics, Laplacian variance and the
// HUD temporal-variance mask.  Integer wherever the reference is integer (exact, order-free);
// floating point only in the final rational / mean, reduced in a fixed order (deterministic).
#include <math.h>

#include "common.cuh"

__device__ __forceinline__ int gray_u8(int r, int g, int b) {
  // cv2.cvtColor(RGB2GRAY) integer identity
  return (r * 9798 + g * 19235 + b * 3735 + 16384) >> 15;
}
// cv2.resize(uint8, INTER_LINEAR) restated bit-exactly (OpenCV imgproc resize.cpp: resizeGeneric_ with
// HResizeLinear<uchar,int,short,2048> / VResizeLinear<uchar,int,short,FixedPtCast<int,uchar,22>>; pinned against cv2 4.13 by
// tests/test_oracle_router.py through oracle/resize_ref.py).  Destination index d reads source indices s0, s1 with 11-bit
// coefficients a0 + a1 = 2048.  Along x the fraction is zeroed where the tap pair leaves the row; along y OpenCV keeps the
// fraction and only clips the row indices.  For even integer ratios this is the 2x2 centre average of SURVEY App. B.1.
struct LinTap { int s0, s1, a0, a1; };
__device__ __forceinline__ LinTap lin_tap(int d, double scale, int src, bool clamp_frac) {
  float f = (float)__dsub_rn(__dmul_rn((double)d + 0.5, scale), 0.5);
  int s = (int)floorf(f);
  f = __fsub_rn(f, (float)s);
  if (clamp_frac) {
    if (s < 0) { s = 0; f = 0.f; }
    if (s >= src - 1) { s = src - 1; f = 0.f; }
  }
  LinTap t;
  t.a0 = __float2int_rn(__fmul_rn(__fsub_rn(1.f, f), 2048.f));
  t.a1 = __float2int_rn(__fmul_rn(f, 2048.f));
  t.s0 = min(max(s, 0), src - 1);
  t.s1 = min(max(s + 1, 0), src - 1);
  return t;
}
__device__ __forceinline__ int gray_at(const uint8_t* __restrict__ rgb, int W, int y, int x) {
  const uint8_t* q = rgb + ((size_t)y * W + x) * 3;
  return gray_u8(q[0], q[1], q[2]);
}
// one pixel of cv2.resize(cv2.cvtColor(rgb, RGB2GRAY), ...)
__device__ __forceinline__ int resized_gray(const uint8_t* __restrict__ rgb, int W, const LinTap tx, const LinTap ty) {
  const int h0 = gray_at(rgb, W, ty.s0, tx.s0) * tx.a0 + gray_at(rgb, W, ty.s0, tx.s1) * tx.a1;
  const int h1 = gray_at(rgb, W, ty.s1, tx.s0) * tx.a0 + gray_at(rgb, W, ty.s1, tx.s1) * tx.a1;
  return (((ty.a0 * (h0 >> 4)) >> 16) + ((ty.a1 * (h1 >> 4)) >> 16) + 2) >> 2;
}
template <typename T>
__device__ __forceinline__ T warp_sum(T v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  return v;
}
template <typename T>
__device__ __forceinline__ T block_sum(T v, T* sh /*>=32*/) {
  v = warp_sum(v);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  __syncthreads();
  if (l == 0) sh[w] = v;
  __syncthreads();
  const int nw = (blockDim.x + 31) >> 5;
  v = (threadIdx.x < nw) ? sh[threadIdx.x] : T(0);
  if (w == 0) v = warp_sum(v);
  return v;  // valid in thread 0
}

// ---------------------------------------------------------------------------------------------
// PSNR: exact integer SSE (evaluation/metrics.py:299-315).  16-byte loads, u32 lane accumulators
// flushed to u64, shuffle + smem block reduction, one integer atomic per block.
#define PSNR_BLOCKS_PER_FRAME 296
__global__ void psnr_sse_kernel(const uint8_t* __restrict__ a, const uint8_t* __restrict__ b, size_t elems,
                                unsigned long long* __restrict__ sse) {
  __shared__ unsigned long long sh[32];
  const int f = blockIdx.y;
  const uint8_t* pa = a + (size_t)f * elems;
  const uint8_t* pb = b + (size_t)f * elems;
  unsigned long long acc = 0;
  const bool aligned = ((((uintptr_t)pa) | ((uintptr_t)pb)) & 15) == 0;
  size_t nvec = aligned ? elems / 16 : 0;
  const uint4* va = reinterpret_cast<const uint4*>(pa);
  const uint4* vb = reinterpret_cast<const uint4*>(pb);
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < nvec; i += (size_t)gridDim.x * blockDim.x) {
    uint4 x = __ldg(va + i), y = __ldg(vb + i);
    const uint32_t xs[4] = {x.x, x.y, x.z, x.w}, ys[4] = {y.x, y.y, y.z, y.w};
    uint32_t s = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) {
#pragma unroll
      for (int j = 0; j < 4; j++) {
        int d = (int)((xs[k] >> (8 * j)) & 0xff) - (int)((ys[k] >> (8 * j)) & 0xff);
        s += (uint32_t)(d * d);
      }
    }
    acc += s;
  }
  for (size_t i = nvec * 16 + blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < elems;
       i += (size_t)gridDim.x * blockDim.x) {
    int d = (int)pa[i] - (int)pb[i];
    acc += (unsigned long long)(d * d);
  }
  acc = block_sum(acc, sh);
  if (threadIdx.x == 0 && acc) atomicAdd(sse + f, acc);
}
__global__ void psnr_finish_kernel(const unsigned long long* sse, size_t elems, int n, unsigned long long* sse_out,
                                   double* psnr_out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  unsigned long long s = sse[i];
  if (sse_out) sse_out[i] = s;
  if (psnr_out) {
    double mse = (double)s / (double)elems;
    psnr_out[i] = s == 0 ? INFINITY : 10.0 * log10(255.0 * 255.0 / mse);
  }
}
extern "C" __attribute__((visibility("default"))) int vfi_psnr_u8(vfi_handle* h, const uint8_t* a, const uint8_t* b, int n, size_t elems,
                           unsigned long long* sse_out, double* psnr_out, void* stream) {
  if (!h) return VFI_E_ARG;
  if (!a || !b || n < 1 || elems < 1 || n > 65535) VFI_FAIL(h, VFI_E_ARG, "bad argument");
  VFI_ENTER(h, stream);
  cudaStream_t s = (cudaStream_t)stream;
  if (int rc = vfi_red_reserve(h, (size_t)65536 * 8)) return rc;
  unsigned long long* acc = (unsigned long long*)h->red;
  VFI_CUDA(h, cudaMemsetAsync(acc, 0, (size_t)n * 8, s));
  int bx = (int)std::min<size_t>(PSNR_BLOCKS_PER_FRAME, (elems / 16 + 255) / 256 + 1);
  if (n >= 64) bx = std::min(bx, 32);
  psnr_sse_kernel<<<dim3(bx, n), 256, 0, s>>>(a, b, elems, acc);
  VFI_LAUNCHED(h);
  psnr_finish_kernel<<<cdiv(n, 128), 128, 0, s>>>(acc, elems, n, sse_out, psnr_out);
  VFI_LAUNCHED(h);
  return VFI_OK;
}

// ---------------------------------------------------------------------------------------------
// SSIM (skimage structural_similarity, 2-D uint8, win 7 uniform, sample covariance, data_range 255;
// SURVEY App. B.2) on the cv2 gray image (DS = 1: evaluation/metrics.py:318-335) or on its x0.25
// INTER_LINEAR thumbnail (DS = 4: adaptive_pipeline.py:149-158, = 2x2 centre average, App. B.1).
// Only the valid (h-6)x(w-6) windows contribute (skimage crops the 3-px border), so the window sums
// are exact int32.  Tile 32x32 outputs <- 38x38 gray pixels in smem; separable 7-tap box sums.
#define SS_T 32
#define SS_IN (SS_T + 6)
template <int DS>
__global__ void ssim_kernel(const uint8_t* __restrict__ a, const uint8_t* __restrict__ b, int H, int W, int h, int w,
                            double* __restrict__ partial) {
  __shared__ uint8_t ga[SS_IN][SS_IN + 2], gb[SS_IN][SS_IN + 2];
  __shared__ int hs[5][SS_IN][SS_T + 1];
  __shared__ double shd[32];
  const int f = blockIdx.z;
  const uint8_t* pa = a + (size_t)f * H * W * 3;
  const uint8_t* pb = b + (size_t)f * H * W * 3;
  const int ox = blockIdx.x * SS_T, oy = blockIdx.y * SS_T;
  for (int i = threadIdx.x; i < SS_IN * SS_IN; i += blockDim.x) {
    int ly = i / SS_IN, lx = i % SS_IN;
    int gy = oy + ly, gx = ox + lx;
    int va = 0, vb = 0;
    if (gy < h && gx < w) {
      if (DS == 1) {
        const uint8_t* q = pa + ((size_t)gy * W + gx) * 3;
        va = gray_u8(q[0], q[1], q[2]);
        q = pb + ((size_t)gy * W + gx) * 3;
        vb = gray_u8(q[0], q[1], q[2]);
      } else {  // cv2.resize(gray, None, fx=0.25, fy=0.25): scale = 1 / 0.25
        const LinTap tx = lin_tap(gx, (double)DS, W, true), ty = lin_tap(gy, (double)DS, H, false);
        va = resized_gray(pa, W, tx, ty);
        vb = resized_gray(pb, W, tx, ty);
      }
    }
    ga[ly][lx] = (uint8_t)va; gb[ly][lx] = (uint8_t)vb;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < SS_IN * SS_T; i += blockDim.x) {
    int ly = i / SS_T, lx = i % SS_T;
    int s0 = 0, s1 = 0, s2 = 0, s3 = 0, s4 = 0;
#pragma unroll
    for (int k = 0; k < 7; k++) {
      int x = ga[ly][lx + k], y = gb[ly][lx + k];
      s0 += x; s1 += y; s2 += x * x; s3 += y * y; s4 += x * y;
    }
    hs[0][ly][lx] = s0; hs[1][ly][lx] = s1; hs[2][ly][lx] = s2; hs[3][ly][lx] = s3; hs[4][ly][lx] = s4;
  }
  __syncthreads();
  double local = 0.0;
  for (int i = threadIdx.x; i < SS_T * SS_T; i += blockDim.x) {
    int ly = i / SS_T, lx = i % SS_T;
    if (oy + ly + 6 < h && ox + lx + 6 < w) {
      int s0 = 0, s1 = 0, s2 = 0, s3 = 0, s4 = 0;
#pragma unroll
      for (int k = 0; k < 7; k++) {
        s0 += hs[0][ly + k][lx]; s1 += hs[1][ly + k][lx]; s2 += hs[2][ly + k][lx]; s3 += hs[3][ly + k][lx];
        s4 += hs[4][ly + k][lx];
      }
      const double n = 49.0, cn = 49.0 / 48.0;
      double ux = s0 / n, uy = s1 / n;
      double vx = cn * (s2 / n - ux * ux), vy = cn * (s3 / n - uy * uy), vxy = cn * (s4 / n - ux * uy);
      const double C1 = 6.5025, C2 = 58.5225;
      local += ((2 * ux * uy + C1) * (2 * vxy + C2)) / ((ux * ux + uy * uy + C1) * (vx + vy + C2));
    }
  }
  local = block_sum(local, shd);
  if (threadIdx.x == 0)
    partial[((size_t)f * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x] = local;
}
__global__ void sum_partials_kernel(const double* __restrict__ partial, int per, double denom, double* __restrict__ out) {
  __shared__ double shd[32];
  const int f = blockIdx.x;
  double s = 0.0;
  for (int i = threadIdx.x; i < per; i += blockDim.x) s += partial[(size_t)f * per + i];
  s = block_sum(s, shd);
  if (threadIdx.x == 0) out[f] = s / denom;
}
extern "C" __attribute__((visibility("default"))) int vfi_ssim_gray_u8(vfi_handle* h, const uint8_t* a, const uint8_t* b, int n, int H, int W, int downscale,
                                double* out, void* stream) {
  if (!h) return VFI_E_ARG;
  if (!a || !b || !out || n < 1 || n > 65535) VFI_FAIL(h, VFI_E_ARG, "bad argument");
  if (downscale != 1 && downscale != 4) VFI_FAIL(h, VFI_E_ARG, "downscale must be 1 or 4");
  // cv2.resize(fx=fy=0.25): dsize = cvRound(size * 0.25), round-half-to-even
  const int hh = downscale == 1 ? H : (int)nearbyint(H * 0.25), ww = downscale == 1 ? W : (int)nearbyint(W * 0.25);
  if (hh < 7 || ww < 7) VFI_FAIL(h, VFI_E_ARG, "image smaller than the 7x7 SSIM window");
  VFI_ENTER(h, stream);
  cudaStream_t s = (cudaStream_t)stream;
  dim3 grid(cdiv(ww - 6, SS_T), cdiv(hh - 6, SS_T), n);
  const int per = grid.x * grid.y;
  if (int rc = vfi_red_reserve(h, std::max((size_t)65536 * 8, (size_t)n * per * 8))) return rc;
  double* partial = (double*)h->red;
  if (downscale == 1) ssim_kernel<1><<<grid, 256, 0, s>>>(a, b, H, W, hh, ww, partial);
  else ssim_kernel<4><<<grid, 256, 0, s>>>(a, b, H, W, hh, ww, partial);
  VFI_LAUNCHED(h);
  sum_partials_kernel<<<n, 256, 0, s>>>(partial, per, (double)(hh - 6) * (double)(ww - 6), out);
  VFI_LAUNCHED(h);
  return VFI_OK;
}

// ---------------------------------------------------------------------------------------------
// Flow statistics (adaptive_pipeline.py:130-138): mag = sqrt(u^2+v^2) in fp32 (as numpy does on the
// float32 flow), then mean / max / population std.  fp64 accumulators, fixed-order two-stage reduce.
#define FS_BLOCKS 296
__global__ void flow_stats_kernel(const float* __restrict__ flow, size_t npix, double* __restrict__ partial) {
  __shared__ double shd[32];
  const int f = blockIdx.y;
  const float2* p = reinterpret_cast<const float2*>(flow) + (size_t)f * npix;
  double s = 0.0, s2 = 0.0;
  float mx = 0.f;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < npix; i += (size_t)gridDim.x * blockDim.x) {
    float2 v = __ldg(p + i);
    float m = sqrtf(__fadd_rn(__fmul_rn(v.x, v.x), __fmul_rn(v.y, v.y)));
    s += (double)m;
    s2 += (double)m * (double)m;
    mx = fmaxf(mx, m);
  }
  s = block_sum(s, shd);
  s2 = block_sum(s2, shd);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_down_sync(0xffffffffu, mx, o));
  __shared__ float shm[32];
  __syncthreads();
  if ((threadIdx.x & 31) == 0) shm[threadIdx.x >> 5] = mx;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int i = 1; i < (int)(blockDim.x >> 5); i++) mx = fmaxf(mx, shm[i]);
    double* o = partial + ((size_t)f * gridDim.x + blockIdx.x) * 3;
    o[0] = s; o[1] = s2; o[2] = (double)mx;
  }
}
__global__ void flow_stats_finish_kernel(const double* __restrict__ partial, int per, double npix, double* __restrict__ out) {
  __shared__ double shd[32];
  const int f = blockIdx.x;
  double s = 0.0, s2 = 0.0, mx = 0.0;
  for (int i = threadIdx.x; i < per; i += blockDim.x) {
    const double* p = partial + ((size_t)f * per + i) * 3;
    s += p[0]; s2 += p[1]; mx = fmax(mx, p[2]);
  }
  s = block_sum(s, shd);
  s2 = block_sum(s2, shd);
  __shared__ double shx[256];
  shx[threadIdx.x] = mx;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int i = 1; i < (int)blockDim.x; i++) mx = fmax(mx, shx[i]);
    double mean = s / npix;
    double var = s2 / npix - mean * mean;
    out[3 * f] = mean; out[3 * f + 1] = mx; out[3 * f + 2] = sqrt(var > 0.0 ? var : 0.0);
  }
}
extern "C" __attribute__((visibility("default"))) int vfi_flow_stats(vfi_handle* h, const float* flow, int n, int H, int W, double* out, void* stream) {
  if (!h) return VFI_E_ARG;
  if (!flow || !out || n < 1 || n > 4096 || H < 1 || W < 1) VFI_FAIL(h, VFI_E_ARG, "bad argument");
  VFI_ENTER(h, stream);
  return vfi_flow_stats_impl(h, flow, n, H, W, out, (cudaStream_t)stream);
}
int vfi_flow_stats_impl(vfi_handle* h, const float* flow, int n, int H, int W, double* out, cudaStream_t s) {
  const size_t npix = (size_t)H * W;
  int bx = (int)std::min<size_t>(FS_BLOCKS, (npix + 255) / 256);
  if (int rc = vfi_red_reserve(h, std::max((size_t)65536 * 8, (size_t)n * bx * 24))) return rc;
  double* partial = (double*)h->red;
  flow_stats_kernel<<<dim3(bx, n), 256, 0, s>>>(flow, npix, partial);
  VFI_LAUNCHED(h);
  flow_stats_finish_kernel<<<n, 256, 0, s>>>(partial, bx, (double)npix, out);
  VFI_LAUNCHED(h);
  return VFI_OK;
}

// ---------------------------------------------------------------------------------------------
// Laplacian variance (adaptive_pipeline.py:180-183): cv2.Laplacian(gray, CV_64F), ksize 1 =
// [0 1 0; 1 -4 1; 0 1 0] with BORDER_REFLECT_101, then .var().  The stencil is integer, so
// sum(x) and sum(x^2) are exact int64; only the final variance is fp64.
#define LP_T 32
__global__ void laplacian_kernel(const uint8_t* __restrict__ rgb, int H, int W, long long* __restrict__ acc) {
  __shared__ int g[LP_T + 2][LP_T + 2];
  __shared__ long long shl[32];
  const int f = blockIdx.z;
  const uint8_t* p = rgb + (size_t)f * H * W * 3;
  const int ox = blockIdx.x * LP_T, oy = blockIdx.y * LP_T;
  for (int i = threadIdx.x; i < (LP_T + 2) * (LP_T + 2); i += blockDim.x) {
    int ly = i / (LP_T + 2), lx = i % (LP_T + 2);
    int gy = oy + ly - 1, gx = ox + lx - 1;
    if (gy < 0) gy = -gy;                   // reflect-101
    if (gx < 0) gx = -gx;
    if (gy >= H) gy = 2 * (H - 1) - gy;
    if (gx >= W) gx = 2 * (W - 1) - gx;
    gy = max(0, min(gy, H - 1)); gx = max(0, min(gx, W - 1));
    const uint8_t* q = p + ((size_t)gy * W + gx) * 3;
    g[ly][lx] = gray_u8(q[0], q[1], q[2]);
  }
  __syncthreads();
  long long s = 0, s2 = 0;
  for (int i = threadIdx.x; i < LP_T * LP_T; i += blockDim.x) {
    int ly = i / LP_T, lx = i % LP_T;
    if (oy + ly < H && ox + lx < W) {
      int v = g[ly][lx + 1] + g[ly + 2][lx + 1] + g[ly + 1][lx] + g[ly + 1][lx + 2] - 4 * g[ly + 1][lx + 1];
      s += v; s2 += (long long)v * v;
    }
  }
  s = block_sum(s, shl);
  s2 = block_sum(s2, shl);
  if (threadIdx.x == 0) {
    atomicAdd((unsigned long long*)(acc + 2 * f), (unsigned long long)s);
    atomicAdd((unsigned long long*)(acc + 2 * f + 1), (unsigned long long)s2);
  }
}
__global__ void laplacian_finish_kernel(const long long* acc, int n, double npix, double* out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double mean = (double)acc[2 * i] / npix;
  out[i] = (double)acc[2 * i + 1] / npix - mean * mean;
}
extern "C" __attribute__((visibility("default"))) int vfi_laplacian_var(vfi_handle* h, const uint8_t* rgb, int n, int H, int W, double* out, void* stream) {
  if (!h) return VFI_E_ARG;
  if (!rgb || !out || n < 1 || n > 4096 || H < 2 || W < 2) VFI_FAIL(h, VFI_E_ARG, "bad argument");
  VFI_ENTER(h, stream);
  cudaStream_t s = (cudaStream_t)stream;
  if (int rc = vfi_red_reserve(h, (size_t)65536 * 8)) return rc;
  long long* acc = (long long*)h->red;
  VFI_CUDA(h, cudaMemsetAsync(acc, 0, (size_t)n * 16, s));
  laplacian_kernel<<<dim3(cdiv(W, LP_T), cdiv(H, LP_T), n), 256, 0, s>>>(rgb, H, W, acc);
  VFI_LAUNCHED(h);
  laplacian_finish_kernel<<<cdiv(n, 128), 128, 0, s>>>(acc, n, (double)H * W, out);
  VFI_LAUNCHED(h);
  return VFI_OK;
}

// ---------------------------------------------------------------------------------------------
// HUD mask (adaptive_pipeline.py:193-235).
//   thumb  : gray -> cv2.resize(320x180) INTER_LINEAR = 2x2 centre average for even integer ratios
//   var    : np.var(last 5 thumbs, axis=0) < thr, restated with numpy's exact operation order
//            (mean = sum/5; sum of (x-mean)^2 oldest->newest; /5) so the threshold flag is bit-exact
//   up     : cv2 INTER_NEAREST (src = min(floor(dst * (1/fx)), n-1), double arithmetic as OpenCV)
//   morph  : MORPH_CLOSE then MORPH_OPEN with 5x5 ones = dilate5, erode9, dilate5 (borders ignored)
#define HUD_W 320
#define HUD_H 180
__global__ void hud_thumb_kernel(const uint8_t* __restrict__ rgb, int H, int W, double sy, double sx, uint8_t* __restrict__ slot) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= HUD_W * HUD_H) return;
  int x = i % HUD_W, y = i / HUD_W;
  slot[i] = (uint8_t)resized_gray(rgb, W, lin_tap(x, sx, W, true), lin_tap(y, sy, H, false));
}
struct HudSlots { int idx[5]; };
__global__ void hud_var_kernel(const uint8_t* __restrict__ ring, HudSlots sl, double thr, uint8_t* __restrict__ small) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= HUD_W * HUD_H) return;
  double x[5];
  double sum = 0.0;
#pragma unroll
  for (int k = 0; k < 5; k++) { x[k] = (double)ring[(size_t)sl.idx[k] * HUD_W * HUD_H + i]; sum = __dadd_rn(sum, x[k]); }
  double mean = __ddiv_rn(sum, 5.0);
  double acc = 0.0;
#pragma unroll
  for (int k = 0; k < 5; k++) { double d = __dsub_rn(x[k], mean); acc = __dadd_rn(acc, __dmul_rn(d, d)); }
  double var = __ddiv_rn(acc, 5.0);
  small[i] = var < thr ? 1 : 0;
}
// var flag -> NEAREST upsample -> close -> open -> count, in one kernel.  A CTA owns a 32 x 32 tile of the mask and works on
// the 48 x 48 region around it (the three morphology steps reach 2 + 4 + 2 pixels) as 48 row bit-words: a horizontal
// step is shifts and ORs of one word, a vertical step ORs neighbouring rows, erosion is dilation of the complement, and
// "out-of-image samples are ignored" is a mask on each step's input.
#define HM_T 32
#define HM_R 8
#define HM_N (HM_T + 2 * HM_R)
#define HM_CELLS 50  // small-mask cells a 48-pixel span can touch when the frame is at least 320 x 180 (ratio >= 1);
                     // smaller frames (nearest UPsampling of the thumbnail grid) take the unstaged path
__device__ __forceinline__ bool hud_flag_at(const uint8_t* __restrict__ ring, const int* idx, double thr, int i) {
  double x[5];
  double sum = 0.0;
#pragma unroll
  for (int k = 0; k < 5; k++) { x[k] = (double)ring[(size_t)idx[k] * HUD_W * HUD_H + i]; sum = __dadd_rn(sum, x[k]); }
  const double mean = __ddiv_rn(sum, 5.0);
  double acc = 0.0;
#pragma unroll
  for (int k = 0; k < 5; k++) { const double d = __dsub_rn(x[k], mean); acc = __dadd_rn(acc, __dmul_rn(d, d)); }
  return __ddiv_rn(acc, 5.0) < thr;
}
__device__ __forceinline__ void hm_dilate(unsigned long long* a, unsigned long long* tmp, const bool rowin, unsigned long long inx,
                                          int r, bool complement) {
  const int t = threadIdx.x;
  if (t < HM_N) {
    unsigned long long w = a[t];
    if (complement) w = ~w;
    w = rowin ? (w & inx) : 0ull;
    unsigned long long hz = w;
    for (int k = 1; k <= r; k++) hz |= (w << k) | (w >> k);
    tmp[t] = hz;
  }
  __syncthreads();
  if (t < HM_N) {
    unsigned long long v = 0ull;
    for (int k = -r; k <= r; k++)
      if (t + k >= 0 && t + k < HM_N) v |= tmp[t + k];
    a[t] = complement ? ~v : v;
  }
  __syncthreads();
}
__global__ void __launch_bounds__(256) hud_mask_kernel(const uint8_t* __restrict__ ring, HudSlots sl, double thr,
                                                       uint8_t* __restrict__ dst, int H, int W, double ify, double ifx,
                                                       unsigned long long* __restrict__ cnt) {
  __shared__ unsigned long long rows[HM_N], tmp[HM_N];
  __shared__ uint8_t flag[HM_CELLS][HM_CELLS];
  const int tx0 = blockIdx.x * HM_T, ty0 = blockIdx.y * HM_T;
  const int ox = tx0 - HM_R, oy = ty0 - HM_R;
  const int x_lo = max(ox, 0), x_hi = min(ox + HM_N - 1, W - 1), y_lo = max(oy, 0), y_hi = min(oy + HM_N - 1, H - 1);
  const int cx0 = min((int)floor(__dmul_rn((double)x_lo, ifx)), HUD_W - 1), cx1 = min((int)floor(__dmul_rn((double)x_hi, ifx)), HUD_W - 1);
  const int cy0 = min((int)floor(__dmul_rn((double)y_lo, ify)), HUD_H - 1), cy1 = min((int)floor(__dmul_rn((double)y_hi, ify)), HUD_H - 1);
  const int ncx = cx1 - cx0 + 1, ncy = cy1 - cy0 + 1;
  const bool staged = ncx <= HM_CELLS && ncy <= HM_CELLS;
  if (staged)
    for (int e = threadIdx.x; e < ncx * ncy; e += blockDim.x) {
      const int cy = e / ncx, cx = e - cy * ncx;
      flag[cy][cx] = hud_flag_at(ring, sl.idx, thr, (cy0 + cy) * HUD_W + cx0 + cx) ? 1 : 0;
    }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int t = warp; t < HM_N; t += 8) {
    const int gy = oy + t;
    const bool yin = gy >= 0 && gy < H;
    const int cya = yin ? min((int)floor(__dmul_rn((double)gy, ify)), HUD_H - 1) : 0, cy = cya - cy0;
    unsigned int half[2];
#pragma unroll
    for (int hh = 0; hh < 2; hh++) {
      const int c = lane + 32 * hh, gx = ox + c;
      bool bit = false;
      if (yin && c < HM_N && gx >= 0 && gx < W) {
        const int cxa = min((int)floor(__dmul_rn((double)gx, ifx)), HUD_W - 1);
        bit = staged ? flag[cy][cxa - cx0] != 0 : hud_flag_at(ring, sl.idx, thr, cya * HUD_W + cxa);
      }
      half[hh] = __ballot_sync(0xffffffffu, bit);
    }
    if (lane == 0) rows[t] = (unsigned long long)half[0] | ((unsigned long long)half[1] << 32);
  }
  __syncthreads();
  // in-image columns of the region as a bit mask; this thread's row
  unsigned long long inx = 0ull;
  {
    const int c_lo = x_lo - ox, c_hi = x_hi - ox;  // 0 <= c_lo <= c_hi < 48
    inx = ((c_hi == 63 ? 0ull : (1ull << (c_hi + 1))) - 1ull) & ~((1ull << c_lo) - 1ull);
  }
  const int gy_t = oy + (int)threadIdx.x;
  const bool rowin = threadIdx.x < HM_N && gy_t >= 0 && gy_t < H;
  hm_dilate(rows, tmp, rowin, inx, 2, false);  // MORPH_CLOSE: dilate 5x5 ...
  hm_dilate(rows, tmp, rowin, inx, 4, true);   // ... erode 5x5, then MORPH_OPEN's erode 5x5: together erode 9x9
  hm_dilate(rows, tmp, rowin, inx, 2, false);  // ... dilate 5x5
  // store the tile (4 pixels per thread) and count it
  const int yl = threadIdx.x >> 3, xq = (threadIdx.x & 7) * 4;
  const int y = ty0 + yl, x = tx0 + xq;
  unsigned int c4 = 0;
  if (y < H && x < W) {
    const unsigned long long w = rows[HM_R + yl] >> (HM_R + xq);
    uint8_t v[4];
#pragma unroll
    for (int k = 0; k < 4; k++) { v[k] = (x + k < W) ? (uint8_t)((w >> k) & 1ull) : 0; c4 += v[k]; }
    uint8_t* o = dst + (size_t)y * W + x;
    if (x + 3 < W && ((W & 3) == 0)) *reinterpret_cast<uchar4*>(o) = make_uchar4(v[0], v[1], v[2], v[3]);
    else for (int k = 0; k < 4 && x + k < W; k++) o[k] = v[k];
  }
  __shared__ unsigned long long sh[32];
  const unsigned long long c = block_sum((unsigned long long)c4, sh);
  if (threadIdx.x == 0 && c) atomicAdd(cnt, c);
}
__global__ void coverage_finish_kernel(const unsigned long long* cnt, double n, double* out) { *out = (double)(*cnt) / n; }

extern "C" __attribute__((visibility("default"))) int vfi_hud_update(vfi_handle* h, const uint8_t* rgb, int H, int W, uint8_t* hist_ring, int hist_len,
                              long long* count, double var_threshold, uint8_t* mask_out, double* coverage_out,
                              void* stream) {
  if (!h) return VFI_E_ARG;
  if (!rgb || !hist_ring || !count || !mask_out || !coverage_out || hist_len < 5) VFI_FAIL(h, VFI_E_ARG, "bad argument");
  if (H < 1 || W < 1) VFI_FAIL(h, VFI_E_ARG, "bad frame size %dx%d", H, W);
  VFI_ENTER(h, stream);
  cudaStream_t s = (cudaStream_t)stream;
  const size_t npix = (size_t)H * W;
  if (int rc = vfi_red_reserve(h, std::max((size_t)65536 * 8, npix + HUD_W * HUD_H + 4096))) return rc;
  const long long c = *count;
  // cv2.resize(gray, (320, 180)): inv_scale = dsize / ssize, scale = 1 / inv_scale (doubles, as OpenCV computes them)
  const double rsx = 1.0 / ((double)HUD_W / (double)W), rsy = 1.0 / ((double)HUD_H / (double)H);
  hud_thumb_kernel<<<cdiv(HUD_W * HUD_H, 256), 256, 0, s>>>(rgb, H, W, rsy, rsx,
                                                            hist_ring + (size_t)(c % hist_len) * HUD_W * HUD_H);
  VFI_LAUNCHED(h);
  *count = c + 1;
  if (c + 1 < 5) {
    VFI_CUDA(h, cudaMemsetAsync(mask_out, 0, npix, s));
    VFI_CUDA(h, cudaMemsetAsync(coverage_out, 0, sizeof(double), s));
    return VFI_OK;
  }
  unsigned long long* cnt = (unsigned long long*)h->red;
  HudSlots sl;
  for (int k = 0; k < 5; k++) sl.idx[k] = (int)((c - 4 + k) % hist_len);  // oldest -> newest of the last five
  const double ifx = 1.0 / ((double)W / (double)HUD_W), ify = 1.0 / ((double)H / (double)HUD_H);
  VFI_CUDA(h, cudaMemsetAsync(cnt, 0, 8, s));
  hud_mask_kernel<<<dim3(cdiv(W, HM_T), cdiv(H, HM_T)), 256, 0, s>>>(hist_ring, sl, var_threshold, mask_out, H, W, ify, ifx, cnt);
  VFI_LAUNCHED(h);
  coverage_finish_kernel<<<1, 1, 0, s>>>(cnt, (double)npix, coverage_out);
  VFI_LAUNCHED(h);
  return VFI_OK;
}
