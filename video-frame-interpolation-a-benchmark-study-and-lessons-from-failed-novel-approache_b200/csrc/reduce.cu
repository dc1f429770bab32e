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
  ProfScope ps(h, PC_REDUCE, s);
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
// Global (single-window) moments of a uint8 pair: sum a, sum b, sum a^2, sum b^2, sum ab -- the integers behind
// scripts/run_experiments.py:292-312 calc_psnr / calc_ssim.  16-byte loads, five dp4a per 4 bytes, u64 lane
// accumulators, shuffle + smem block reduction, one integer atomic per block and moment.
__global__ void moments_kernel(const uint8_t* __restrict__ a, const uint8_t* __restrict__ b, size_t elems,
                               unsigned long long* __restrict__ acc5) {
  __shared__ unsigned long long sh[32];
  const int f = blockIdx.y;
  const uint8_t* pa = a + (size_t)f * elems;
  const uint8_t* pb = b + (size_t)f * elems;
  unsigned long long m[5] = {0, 0, 0, 0, 0};
  const bool aligned = ((((uintptr_t)pa) | ((uintptr_t)pb)) & 15) == 0;
  const size_t nvec = aligned ? elems / 16 : 0;
  const uint4* va = reinterpret_cast<const uint4*>(pa);
  const uint4* vb = reinterpret_cast<const uint4*>(pb);
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < nvec; i += (size_t)gridDim.x * blockDim.x) {
    const uint4 x = __ldg(va + i), y = __ldg(vb + i);
    const uint32_t xs[4] = {x.x, x.y, x.z, x.w}, ys[4] = {y.x, y.y, y.z, y.w};
    uint32_t s0 = 0, s1 = 0, s2 = 0, s3 = 0, s4 = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) {
      s0 = __dp4a(xs[k], 0x01010101u, s0); s1 = __dp4a(ys[k], 0x01010101u, s1);
      s2 = __dp4a(xs[k], xs[k], s2); s3 = __dp4a(ys[k], ys[k], s3); s4 = __dp4a(xs[k], ys[k], s4);
    }
    m[0] += s0; m[1] += s1; m[2] += s2; m[3] += s3; m[4] += s4;
  }
  for (size_t i = nvec * 16 + blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < elems; i += (size_t)gridDim.x * blockDim.x) {
    const unsigned long long x = pa[i], y = pb[i];
    m[0] += x; m[1] += y; m[2] += x * x; m[3] += y * y; m[4] += x * y;
  }
#pragma unroll
  for (int k = 0; k < 5; k++) {
    const unsigned long long v = block_sum(m[k], sh);
    if (threadIdx.x == 0 && v) atomicAdd(acc5 + 5 * f + k, v);
  }
}
extern "C" __attribute__((visibility("default"))) int vfi_moments_u8(vfi_handle* h, const uint8_t* a, const uint8_t* b, int n, size_t elems,
                                                                       long long* out, void* stream) {
  if (!h) return VFI_E_ARG;
  if (!a || !b || !out || n < 1 || elems < 1 || n > 65535) VFI_FAIL(h, VFI_E_ARG, "bad argument");
  if (elems > ((size_t)1 << 46)) VFI_FAIL(h, VFI_E_ARG, "too many elements for exact int64 moments");
  VFI_ENTER(h, stream);
  cudaStream_t s = (cudaStream_t)stream;
  VFI_CUDA(h, cudaMemsetAsync(out, 0, (size_t)n * 5 * 8, s));
  int bx = (int)std::min<size_t>(PSNR_BLOCKS_PER_FRAME, (elems / 16 + 255) / 256 + 1);
  if (n >= 64) bx = std::min(bx, 32);
  ProfScope ps(h, PC_REDUCE, s);
  moments_kernel<<<dim3(bx, n), 256, 0, s>>>(a, b, elems, reinterpret_cast<unsigned long long*>(out));
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
// ---------------------------------------------------------------------------------------------
// Full-resolution SSIM (+ the PSNR sum of squared errors of the same frame pair from the same read): the config-5 kernel.
// A CTA owns a strip of 128 input columns (122 output windows) and a segment of rows and walks DOWN it:
//   V phase  thread = column: gray of both images for the new row (cv2 integers), vertical 7-row running sums of
//            (a | b << 16), a^2 + b^2, a*b -- the old row's pixel comes back from a 7-deep shared-memory ring -- written
//            to a 16-row shared-memory tile;
//   H phase  thread = (row, 16-column run): horizontal 7-column running sums over that tile, the SSIM rational from the
//            five exact integer window sums, in fp64 (one multiply by a correctly rounded reciprocal per window).
// Each input byte is read once (+ 6/128 column and 6/rows halo), 16 tile rows per pair of block barriers.
// SSIM only needs sum(a), sum(b), sum(a^2) + sum(b^2) and sum(ab):
//   S = (2AB/n^2 + C1)(2 cov + C2) / ((A^2 + B^2)/n^2 + C1)(var_a + var_b + C2), n = 49, cov_norm = n/(n-1)
//     = (2AB + C1 n^2)(2(n sum(ab) - AB) + C2 n(n-1)) / (A^2 + B^2 + C1 n^2)(n(sum(a^2) + sum(b^2)) - A^2 - B^2 + C2 n(n-1))
#define SF_COLS 128
#define SF_OUT (SF_COLS - 6)
#define SF_ROWS 16
#define SF_STRIDE (SF_COLS + SF_COLS / 16)  // 16-column runs skewed by one word: the H phase reads conflict-free
__device__ __forceinline__ double ssim_window(int pk, int q, int ab) {
  const int A = pk & 0xffff, B = (unsigned)pk >> 16;
  const int AB = A * B, t3 = A * A + B * B;
  const double c1 = 6.5025 * 2401.0, c2 = 58.5225 * 2352.0;
  const double num = ((double)(2 * AB) + c1) * ((double)(2 * (49 * ab - AB)) + c2);
  const double den = ((double)t3 + c1) * ((double)(49 * q - t3) + c2);
  return num * __drcp_rn(den);
}
template <bool SSE>
__global__ void __launch_bounds__(SF_COLS) ssim_full_kernel(const uint8_t* __restrict__ a, const uint8_t* __restrict__ b, int H, int W,
                                                            int rows_per_seg, double* __restrict__ part_ssim,
                                                            unsigned long long* __restrict__ part_sse) {
  __shared__ int vt[3][SF_ROWS][SF_STRIDE];
  __shared__ uint8_t ring[2][7][SF_COLS];
  __shared__ double shd[32];
  __shared__ unsigned long long shl[32];
  const int t = threadIdx.x, f = blockIdx.z;
  const int x = blockIdx.x * SF_OUT + t;
  const bool xin = x < W;
  const int oh = H - 6;                                           // output rows
  const int y0 = blockIdx.y * rows_per_seg, y1 = min(oh, y0 + rows_per_seg);   // this CTA's output rows
  const bool last_seg = y1 == oh, last_strip = blockIdx.x == gridDim.x - 1;
  const uint8_t* pa = a + ((size_t)f * H * W + x) * 3;
  const uint8_t* pb = b + ((size_t)f * H * W + x) * 3;
  const bool own_col = xin && (t < SF_OUT || last_strip);
  int pk = 0, q = 0, ab = 0;
  unsigned int sse = 0;
  double acc = 0.0;
  const int tpos = t + (t >> 4);
  const int hrow = t >> 3, hseg = t & 7;                          // H-phase role
  const int ow = W - 6;
  int slot = 0;
  for (int r0 = y0; r0 < y1 + 6; r0 += 4) {
    // four rows of loads in flight before any of them is consumed
    int ga[4], gb[4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const int r = r0 + j;
      ga[j] = gb[j] = 0;
      if (xin && r < y1 + 6) {
        const uint8_t* qa = pa + (size_t)r * W * 3;
        const uint8_t* qb = pb + (size_t)r * W * 3;
        const int ar = qa[0], ag = qa[1], abl = qa[2], br = qb[0], bg = qb[1], bb = qb[2];
        ga[j] = gray_u8(ar, ag, abl);
        gb[j] = gray_u8(br, bg, bb);
        if (SSE && own_col && (r < y0 + rows_per_seg || last_seg)) {
          const int d0 = ar - br, d1 = ag - bg, d2 = abl - bb;
          sse += (unsigned)(d0 * d0 + d1 * d1 + d2 * d2);
        }
      }
    }
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const int r = r0 + j;
      if (r >= y1 + 6) break;                                      // uniform over the CTA
      const int k = (r - y0) % 7;
      if (r - y0 >= 7) {                                           // the row leaving the 7-row window
        const int oa = ring[0][k][t], ob = ring[1][k][t];
        pk -= oa | (ob << 16); q -= oa * oa + ob * ob; ab -= oa * ob;
      }
      ring[0][k][t] = (uint8_t)ga[j]; ring[1][k][t] = (uint8_t)gb[j];
      pk += ga[j] | (gb[j] << 16); q += ga[j] * ga[j] + gb[j] * gb[j]; ab += ga[j] * gb[j];
      if (r - y0 >= 6) {                                           // window complete: output row r - 6
        vt[0][slot][tpos] = pk; vt[1][slot][tpos] = q; vt[2][slot][tpos] = ab;
        slot++;
      }
      if (slot == SF_ROWS || (r == y1 + 5 && slot > 0)) {
        __syncthreads();
        if (hrow < slot) {
          const int xl0 = hseg * 16, base = hseg * 17;
          const int nout = min(16, min(SF_OUT - xl0, ow - ((int)blockIdx.x * SF_OUT + xl0)));   // signed: negative = no window
          if (nout > 0) {
            const int* v0 = &vt[0][hrow][base];
            const int* v1 = &vt[1][hrow][base];
            const int* v2 = &vt[2][hrow][base];
            int hp = 0, hq = 0, hab = 0;
#pragma unroll
            for (int c = 0; c < 7; c++) { hp += v0[c]; hq += v1[c]; hab += v2[c]; }
#pragma unroll 4
            for (int i = 0; i < nout; i++) {
              acc += ssim_window(hp, hq, hab);
              const int cn = i + 7 + ((i + 7) >> 4), co = i;      // entering / leaving column (skewed position)
              hp += v0[cn] - v0[co]; hq += v1[cn] - v1[co]; hab += v2[cn] - v2[co];
            }
          }
        }
        __syncthreads();
        slot = 0;
      }
    }
  }
  acc = block_sum(acc, shd);
  const size_t pi = ((size_t)f * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
  if (t == 0) part_ssim[pi] = acc;
  if (SSE) {
    const unsigned long long s = block_sum((unsigned long long)sse, shl);
    if (t == 0) part_sse[pi] = s;
  }
}
// Fast path (W % 4 == 0, 4-byte aligned frames): NO shared memory and NO block barrier.  A warp owns a strip of 128 input
// columns (120 output windows; lane = 4 adjacent columns, read as three aligned 32-bit words per image and row) and a
// segment of rows and walks down it.  Per row and lane: cv2 gray of the 4 + 4 pixels (two dp4a each: the 14-bit
// coefficients split into 7-bit halves), the SSE of the 12 byte pairs (vabsdiffu4 + dp4a), the vertical 7-row running sums of
// (a | b << 16), a^2 + b^2, ab per column -- the leaving row comes from a 7-deep ring of packed grays held in REGISTERS
// (the row loop is unrolled by 7, so the ring index is static) -- then the horizontal 7-column sums from the lane's own
// four columns and six more fetched from lanes +1 / +2 by shuffle, and the SSIM rational of four windows in fp64
// (the integer -> double conversions ride on the fp64 adder through the 2^52 magic constant).
#define SW_OUT 120
__device__ __forceinline__ int gray_dp4a(unsigned px) {   // px = r | g << 8 | b << 16 (byte 3 ignored)
  return (int)(((__dp4a(px, 0x001D964Cu, 0u) << 7) + __dp4a(px, 0x00172346u, 16384u)) >> 15);
}
__device__ __forceinline__ double ssim_window_fast(int pk, int q, int ab) {
  const int A = pk & 0xffff, B = (unsigned)pk >> 16;
  const int AB = A * B, t3 = A * A + B * B;
  const double c1 = 6.5025 * 2401.0, c2 = 58.5225 * 2352.0;
  const double two52 = 4503599627370496.0, two31 = 2147483648.0;
  // as_double(0x43300000 : u) = 2^52 + u exactly, so (that - 2^52) is the exact conversion on the fp64 adder (the I2F
  // unit is 4x slower); the constants are added afterwards: folding them into the 2^52 constant would round them to 0.5
  const double n1 = (__hiloint2double(0x43300000, 2 * AB) - two52) + c1;
  const double n2 = (__hiloint2double(0x43300000, (2 * (49 * ab - AB)) ^ 0x80000000) - (two52 + two31)) + c2;
  const double d1 = (__hiloint2double(0x43300000, t3) - two52) + c1;
  const double d2 = (__hiloint2double(0x43300000, 49 * q - t3) - two52) + c2;
  // 1 / den: MUFU.RCP64H seed (~20 bits) + two Newton steps (relative error ~1e-16); den >= C1 * C2 * n^3 (n - 1) > 0
  const double den = d1 * d2;
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(den));
  r = fma(r, fma(-den, r, 1.0), r);
  r = fma(r, fma(-den, r, 1.0), r);
  return (n1 * n2) * r;
}
template <bool SSE>
__global__ void __launch_bounds__(128, 6) ssim_warp_kernel(const uint8_t* __restrict__ a, const uint8_t* __restrict__ b, int H, int W,
                                                        int rows_per_seg, int strips, double* __restrict__ part_ssim,
                                                        unsigned long long* __restrict__ part_sse) {
  const int lane = threadIdx.x & 31, sx = blockIdx.x * 4 + (threadIdx.x >> 5), f = blockIdx.z;
  if (sx >= strips) return;
  const int x = sx * SW_OUT + lane * 4;
  const bool xin = x < W;
  const int oh = H - 6;
  const int y0 = blockIdx.y * rows_per_seg, y1 = min(oh, y0 + rows_per_seg);
  const bool last_seg = y1 == oh;
  const int yend = y1 + 6;                                  // input rows y0 .. yend-1
  const int wpr = (W * 3) >> 2;                             // 32-bit words per image row
  const uint32_t* pa = reinterpret_cast<const uint32_t*>(a + ((size_t)f * H * W + x) * 3);
  const uint32_t* pb = reinterpret_cast<const uint32_t*>(b + ((size_t)f * H * W + x) * 3);
  const bool own = xin && (lane < SW_OUT / 4 || sx == strips - 1);
  bool valid[4];
#pragma unroll
  for (int c = 0; c < 4; c++) valid[c] = lane < SW_OUT / 4 && x + c + 6 < W;
  int pk[4] = {0, 0, 0, 0}, q[4] = {0, 0, 0, 0}, ab[4] = {0, 0, 0, 0};
  uint32_t ring_a[7] = {0, 0, 0, 0, 0, 0, 0}, ring_b[7] = {0, 0, 0, 0, 0, 0, 0};
  double acc = 0.0;
  unsigned long long sse = 0;
  uint32_t na[3], nb[3];                                    // the NEXT row's words: loads run one row ahead
#pragma unroll
  for (int k = 0; k < 3; k++) {
    na[k] = xin ? __ldg(pa + (size_t)y0 * wpr + k) : 0u;
    nb[k] = xin ? __ldg(pb + (size_t)y0 * wpr + k) : 0u;
  }
  for (int r0 = y0; r0 < yend; r0 += 7) {
#pragma unroll
    for (int j = 0; j < 7; j++) {
      const int r = r0 + j;
      if (r >= yend) break;                                 // warp-uniform
      uint32_t wa[1][3], wb[1][3];
      constexpr int J = 0;
#pragma unroll
      for (int k = 0; k < 3; k++) { wa[0][k] = na[k]; wb[0][k] = nb[k]; }
      {
        const bool ld = xin && r + 1 < yend;
#pragma unroll
        for (int k = 0; k < 3; k++) {
          na[k] = ld ? __ldg(pa + (size_t)(r + 1) * wpr + k) : 0u;
          nb[k] = ld ? __ldg(pb + (size_t)(r + 1) * wpr + k) : 0u;
        }
      }
      if (SSE && own && (r < y0 + rows_per_seg || last_seg)) {
        uint32_t e = 0;
#pragma unroll
        for (int k = 0; k < 3; k++) { const uint32_t d = __vabsdiffu4(wa[J][k], wb[J][k]); e = __dp4a(d, d, e); }
        sse += e;
      }
      const uint32_t pxa[4] = {wa[J][0], __byte_perm(wa[J][0], wa[J][1], 0x0543), __byte_perm(wa[J][1], wa[J][2], 0x0432), wa[J][2] >> 8};
      const uint32_t pxb[4] = {wb[J][0], __byte_perm(wb[J][0], wb[J][1], 0x0543), __byte_perm(wb[J][1], wb[J][2], 0x0432), wb[J][2] >> 8};
      const uint32_t oa4 = ring_a[j], ob4 = ring_b[j];
      uint32_t na4 = 0, nb4 = 0;
#pragma unroll
      for (int c = 0; c < 4; c++) {
        const int ga = gray_dp4a(pxa[c]), gb = gray_dp4a(pxb[c]);
        const int oa = (oa4 >> (8 * c)) & 0xff, ob = (ob4 >> (8 * c)) & 0xff;
        pk[c] += (ga | (gb << 16)) - (oa | (ob << 16));
        q[c] += ga * ga + gb * gb - oa * oa - ob * ob;
        ab[c] += ga * gb - oa * ob;
        na4 |= (uint32_t)ga << (8 * c); nb4 |= (uint32_t)gb << (8 * c);
      }
      ring_a[j] = na4; ring_b[j] = nb4;
      if (r - y0 >= 6) {                                    // 7-row window complete: output row r - 6
        int e0[10], e1[10], e2[10];
#pragma unroll
        for (int c = 0; c < 4; c++) {
          e0[c] = pk[c]; e1[c] = q[c]; e2[c] = ab[c];
          e0[4 + c] = __shfl_down_sync(0xffffffffu, pk[c], 1); e1[4 + c] = __shfl_down_sync(0xffffffffu, q[c], 1);
          e2[4 + c] = __shfl_down_sync(0xffffffffu, ab[c], 1);
        }
#pragma unroll
        for (int c = 0; c < 2; c++) {
          e0[8 + c] = __shfl_down_sync(0xffffffffu, pk[c], 2); e1[8 + c] = __shfl_down_sync(0xffffffffu, q[c], 2);
          e2[8 + c] = __shfl_down_sync(0xffffffffu, ab[c], 2);
        }
        int hp = 0, hq = 0, hab = 0;
#pragma unroll
        for (int i = 0; i < 7; i++) { hp += e0[i]; hq += e1[i]; hab += e2[i]; }
#pragma unroll
        for (int c = 0; c < 4; c++) {
          const double sw = ssim_window_fast(hp, hq, hab);
          acc += valid[c] ? sw : 0.0;                         // select, not a branch (a garbage window may give NaN: dropped)
          if (c < 3) { hp += e0[c + 7] - e0[c]; hq += e1[c + 7] - e1[c]; hab += e2[c + 7] - e2[c]; }
        }
      }
    }
  }
  acc = warp_sum(acc);
  const size_t pi = ((size_t)f * gridDim.y + blockIdx.y) * strips + sx;
  if (lane == 0) part_ssim[pi] = acc;
  if (SSE) {
    sse = warp_sum(sse);
    if (lane == 0) part_sse[pi] = sse;
  }
}
__global__ void ssim_full_finish_kernel(const double* __restrict__ part_ssim, const unsigned long long* __restrict__ part_sse, int per,
                                        double denom, size_t elems, double* __restrict__ ssim_out,
                                        unsigned long long* __restrict__ sse_out, double* __restrict__ psnr_out) {
  __shared__ double shd[32];
  __shared__ unsigned long long shl[32];
  const int f = blockIdx.x;
  double s = 0.0;
  unsigned long long e = 0;
  for (int i = threadIdx.x; i < per; i += blockDim.x) {
    s += part_ssim[(size_t)f * per + i];
    if (part_sse) e += part_sse[(size_t)f * per + i];
  }
  s = block_sum(s, shd);
  if (part_sse) e = block_sum(e, shl);
  if (threadIdx.x == 0) {
    if (ssim_out) ssim_out[f] = s / denom;
    if (sse_out) sse_out[f] = e;
    if (psnr_out) psnr_out[f] = e == 0 ? INFINITY : 10.0 * log10(255.0 * 255.0 / ((double)e / (double)elems));
  }
}
static int ssim_full_launch(vfi_handle* h, const uint8_t* a, const uint8_t* b, int n, int H, int W, bool want_sse,
                            double* ssim_out, unsigned long long* sse_out, double* psnr_out, cudaStream_t s) {
  const int oh = H - 6, ow = W - 6;
  ProfScope ps(h, PC_REDUCE, s);
  if ((W & 3) == 0 && ((((uintptr_t)a) | ((uintptr_t)b)) & 3) == 0) {
    // warp-walking kernel: aim at ~40 warps per SM; at least 48 output rows per segment so the 6-row halo stays cheap
    const int strips = cdiv(ow, SW_OUT);
    int segs = std::max(1, std::min(cdiv(oh, 48), cdiv(40 * h->sm_count, strips * n)));
    const int rps = cdiv(oh, segs);
    segs = cdiv(oh, rps);
    const int per = strips * segs;
    if (int rc = vfi_red_reserve(h, std::max((size_t)65536 * 8, (size_t)n * per * 16 + 64))) return rc;
    double* part = (double*)h->red;
    unsigned long long* parte = (unsigned long long*)(part + (size_t)n * per);
    const dim3 grid(cdiv(strips, 4), segs, n);
    if (want_sse) ssim_warp_kernel<true><<<grid, 128, 0, s>>>(a, b, H, W, rps, strips, part, parte);
    else ssim_warp_kernel<false><<<grid, 128, 0, s>>>(a, b, H, W, rps, strips, part, nullptr);
    VFI_LAUNCHED(h);
    ssim_full_finish_kernel<<<n, 256, 0, s>>>(part, want_sse ? parte : nullptr, per, (double)oh * (double)ow, (size_t)H * W * 3, ssim_out,
                                              sse_out, psnr_out);
    VFI_LAUNCHED(h);
    return VFI_OK;
  }
  const int strips = cdiv(ow, SF_OUT);
  // enough CTAs for ~6 per SM when the batch is small; at least 32 output rows per CTA so the 6-row halo stays cheap
  int segs = std::max(1, std::min(cdiv(oh, 32), cdiv(6 * h->sm_count, strips * n)));
  const int rps = cdiv(oh, segs);
  segs = cdiv(oh, rps);
  const int per = strips * segs;
  if (int rc = vfi_red_reserve(h, std::max((size_t)65536 * 8, (size_t)n * per * 16 + 64))) return rc;
  double* part = (double*)h->red;
  unsigned long long* parte = (unsigned long long*)(part + (size_t)n * per);
  dim3 grid(strips, segs, n);
  if (want_sse) ssim_full_kernel<true><<<grid, SF_COLS, 0, s>>>(a, b, H, W, rps, part, parte);
  else ssim_full_kernel<false><<<grid, SF_COLS, 0, s>>>(a, b, H, W, rps, part, nullptr);
  VFI_LAUNCHED(h);
  ssim_full_finish_kernel<<<n, 256, 0, s>>>(part, want_sse ? parte : nullptr, per, (double)oh * (double)ow, (size_t)H * W * 3, ssim_out,
                                            sse_out, psnr_out);
  VFI_LAUNCHED(h);
  return VFI_OK;
}
extern "C" __attribute__((visibility("default"))) int vfi_psnr_ssim_u8(vfi_handle* h, const uint8_t* a, const uint8_t* b, int n, int H, int W,
                                                                         unsigned long long* sse_out, double* psnr_out, double* ssim_out,
                                                                         void* stream) {
  if (!h) return VFI_E_ARG;
  if (!a || !b || !ssim_out || n < 1 || n > 65535) VFI_FAIL(h, VFI_E_ARG, "bad argument");
  if (H < 7 || W < 7) VFI_FAIL(h, VFI_E_ARG, "image smaller than the 7x7 SSIM window");
  VFI_ENTER(h, stream);
  return ssim_full_launch(h, a, b, n, H, W, true, ssim_out, sse_out, psnr_out, (cudaStream_t)stream);
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
  if (downscale == 1) return ssim_full_launch(h, a, b, n, H, W, false, out, nullptr, nullptr, s);
  ProfScope ps(h, PC_REDUCE, s);
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
  // two pixels per 16-byte load, four loads in flight per thread; the per-pixel arithmetic is numpy's (float32 magnitude)
  const size_t n4 = ((npix & 1) == 0 && (reinterpret_cast<uintptr_t>(p) & 15) == 0) ? npix / 2 : 0;
  const float4* p4 = reinterpret_cast<const float4*>(p);
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i0 = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i0 < n4; i0 += 4 * stride) {
    float4 v[4];
#pragma unroll
    for (int k = 0; k < 4; k++) v[k] = i0 + k * stride < n4 ? __ldg(p4 + i0 + k * stride) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const float m0 = sqrtf(__fadd_rn(__fmul_rn(v[k].x, v[k].x), __fmul_rn(v[k].y, v[k].y)));
      const float m1 = sqrtf(__fadd_rn(__fmul_rn(v[k].z, v[k].z), __fmul_rn(v[k].w, v[k].w)));
      s += (double)m0 + (double)m1;
      s2 += (double)m0 * (double)m0 + (double)m1 * (double)m1;
      mx = fmaxf(mx, fmaxf(m0, m1));
    }
  }
  for (size_t i = 2 * n4 + blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < npix; i += stride) {
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
// A warp owns 30 output columns (lanes 0 and 31 carry the left / right halo column) and walks down its row segment with the
// previous / current / next gray rows in registers: horizontal neighbours come by shuffle, each byte is read once, no
// shared memory and no block barrier in the loop.  Per-lane partial sums stay 32-bit (segments are at most 512 rows).
#define LP_WARPS 4
#define LP_COLS 30
#define LP_MAX_ROWS 512
__device__ __forceinline__ int lp_gray_row(const uint8_t* __restrict__ p, int W, int H, int y, int x, bool xin) {
  if (!xin) return 0;
  y = y < 0 ? -y : (y >= H ? 2 * (H - 1) - y : y);               // reflect-101 (cv2 BORDER_DEFAULT)
  y = max(0, min(y, H - 1));
  const uint8_t* q = p + ((size_t)y * W + x) * 3;
  return gray_u8(q[0], q[1], q[2]);
}
__global__ void __launch_bounds__(LP_WARPS * 32) laplacian_kernel(const uint8_t* __restrict__ rgb, int H, int W, int rows_per_seg,
                                                                 long long* __restrict__ partial) {
  __shared__ long long shl[32];
  const int f = blockIdx.z, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint8_t* p = rgb + (size_t)f * H * W * 3;
  const int xo = (blockIdx.x * LP_WARPS + warp) * LP_COLS + lane - 1;   // lane's own column (halo lanes: one outside the run)
  int xs = xo < 0 ? -xo : (xo >= W ? 2 * (W - 1) - xo : xo);           // reflect-101
  const bool xin = xo <= W && xs >= 0 && xs < W;                        // xo == W is the right halo of the last column
  xs = max(0, min(xs, W - 1));
  const bool owns = lane >= 1 && lane <= LP_COLS && xo < W;
  const int y0 = blockIdx.y * rows_per_seg, y1 = min(H, y0 + rows_per_seg);
  int s = 0;
  unsigned int s2 = 0;
  int gp = lp_gray_row(p, W, H, y0 - 1, xs, xin), gc = lp_gray_row(p, W, H, y0, xs, xin);
  for (int y = y0; y < y1; y += 4) {
    int gn[4];
#pragma unroll
    for (int j = 0; j < 4; j++) gn[j] = lp_gray_row(p, W, H, y + 1 + j, xs, xin && y + j < y1);
#pragma unroll
    for (int j = 0; j < 4; j++) {
      if (y + j >= y1) break;
      const int l = __shfl_up_sync(0xffffffffu, gc, 1), r = __shfl_down_sync(0xffffffffu, gc, 1);
      const int v = gp + gn[j] + l + r - 4 * gc;
      if (owns) { s += v; s2 += (unsigned)(v * v); }
      gp = gc; gc = gn[j];
    }
  }
  const long long bs = block_sum((long long)s, shl);
  const long long bs2 = block_sum((long long)s2, shl);
  if (threadIdx.x == 0) {
    long long* o = partial + 2 * (((size_t)f * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x);
    o[0] = bs; o[1] = bs2;
  }
}
// Fast path (W % 4 == 0, 4-byte aligned frames): a lane owns 4 adjacent columns read as three aligned 32-bit words per row
// (gray by dp4a, as in ssim_warp_kernel); lanes 1..30 of a warp own 120 columns, lanes 0 and 31 carry the neighbouring
// strips' edge columns.  The previous / current / next gray rows live in registers as packed bytes; a row costs three loads,
// four grays and two shuffles per lane.  Reflect-101 needs no extra loads: column -1 is column 1 (the lane's own second
// gray), column W is column W - 2 (its own third), rows -1 / H are rows 1 / H - 2.
#define LW_OUT 120
__device__ __forceinline__ uint32_t lw_gray_row(const uint32_t* __restrict__ p, size_t wpr, int H, int y, bool xin) {
  if (!xin) return 0u;
  y = y < 0 ? -y : (y >= H ? 2 * (H - 1) - y : y);
  y = max(0, min(y, H - 1));
  const uint32_t* q = p + (size_t)y * wpr;
  const uint32_t w0 = __ldg(q), w1 = __ldg(q + 1), w2 = __ldg(q + 2);
  return (uint32_t)gray_dp4a(w0) | ((uint32_t)gray_dp4a(__byte_perm(w0, w1, 0x0543)) << 8) |
         ((uint32_t)gray_dp4a(__byte_perm(w1, w2, 0x0432)) << 16) | ((uint32_t)gray_dp4a(w2 >> 8) << 24);
}
__global__ void __launch_bounds__(128) laplacian_warp_kernel(const uint8_t* __restrict__ rgb, int H, int W, int rows_per_seg, int strips,
                                                             long long* __restrict__ partial) {
  const int lane = threadIdx.x & 31, sx = blockIdx.x * 4 + (threadIdx.x >> 5), f = blockIdx.z;
  if (sx >= strips) return;
  const int x = sx * LW_OUT + (lane - 1) * 4;               // lane 0: the four columns left of the strip
  const bool xin = x >= 0 && x < W;
  const bool owns = xin && lane >= 1 && lane <= LW_OUT / 4;
  const bool first_col = x == 0, last_col = x + 4 == W;
  const size_t wpr = (size_t)(W * 3) >> 2;
  const uint32_t* p = reinterpret_cast<const uint32_t*>(rgb + ((size_t)f * H * W + (xin ? x : 0)) * 3);
  const int y0 = blockIdx.y * rows_per_seg, y1 = min(H, y0 + rows_per_seg);
  int s = 0;
  unsigned long long s2 = 0;
  uint32_t gp = lw_gray_row(p, wpr, H, y0 - 1, xin), gc = lw_gray_row(p, wpr, H, y0, xin);
  for (int y = y0; y < y1; y += 4) {
    uint32_t gn[4];
#pragma unroll
    for (int j = 0; j < 4; j++) gn[j] = lw_gray_row(p, wpr, H, y + 1 + j, xin && y + j < y1);
    uint32_t e2 = 0;
#pragma unroll
    for (int j = 0; j < 4; j++) {
      if (y + j >= y1) break;
      int l = (int)(__shfl_up_sync(0xffffffffu, gc, 1) >> 24), r = (int)(__shfl_down_sync(0xffffffffu, gc, 1) & 0xff);
      if (first_col) l = (gc >> 8) & 0xff;                  // column -1 = column 1
      if (last_col) r = (gc >> 16) & 0xff;                  // column W = column W - 2
      if (owns) {
        const int c0 = gc & 0xff, c1 = (gc >> 8) & 0xff, c2 = (gc >> 16) & 0xff, c3 = gc >> 24;
        const int v0 = (int)(gp & 0xff) + (int)(gn[j] & 0xff) + l + c1 - 4 * c0;
        const int v1 = (int)((gp >> 8) & 0xff) + (int)((gn[j] >> 8) & 0xff) + c0 + c2 - 4 * c1;
        const int v2 = (int)((gp >> 16) & 0xff) + (int)((gn[j] >> 16) & 0xff) + c1 + c3 - 4 * c2;
        const int v3 = (int)(gp >> 24) + (int)(gn[j] >> 24) + c2 + r - 4 * c3;
        s += v0 + v1 + v2 + v3;
        e2 += (uint32_t)(v0 * v0 + v1 * v1 + v2 * v2 + v3 * v3);   // |v| <= 1020: four rows of four stay below 2^32
      }
      gp = gc; gc = gn[j];
    }
    s2 += e2;
  }
  const long long ws = warp_sum((long long)s), ws2 = warp_sum((long long)s2);
  if (lane == 0) {
    long long* o = partial + 2 * (((size_t)f * gridDim.y + blockIdx.y) * strips + sx);
    o[0] = ws; o[1] = ws2;
  }
}
__global__ void laplacian_finish_kernel(const long long* __restrict__ partial, int per, double npix, double* __restrict__ out) {
  __shared__ long long shl[32];
  const int f = blockIdx.x;
  long long s = 0, s2 = 0;
  for (int i = threadIdx.x; i < per; i += blockDim.x) { s += partial[2 * ((size_t)f * per + i)]; s2 += partial[2 * ((size_t)f * per + i) + 1]; }
  s = block_sum(s, shl);
  s2 = block_sum(s2, shl);
  if (threadIdx.x == 0) {
    const double mean = (double)s / npix;
    out[f] = (double)s2 / npix - mean * mean;
  }
}
extern "C" __attribute__((visibility("default"))) int vfi_laplacian_var(vfi_handle* h, const uint8_t* rgb, int n, int H, int W, double* out, void* stream) {
  if (!h) return VFI_E_ARG;
  if (!rgb || !out || n < 1 || n > 4096 || H < 2 || W < 2) VFI_FAIL(h, VFI_E_ARG, "bad argument");
  VFI_ENTER(h, stream);
  cudaStream_t s = (cudaStream_t)stream;
  ProfScope ps(h, PC_REDUCE, s);
  const bool fast = (W & 3) == 0 && W >= 8 && (((uintptr_t)rgb) & 3) == 0;
  const int strips = fast ? cdiv(W, LW_OUT) : cdiv(W, LP_WARPS * LP_COLS);
  // per-lane 32-bit row sums: at most 2^20 rows per segment would overflow nothing, but keep segments short enough to fill the GPU
  int segs = std::max(cdiv(H, LP_MAX_ROWS), std::min(cdiv(H, 16), cdiv((fast ? 32 : 8) * h->sm_count, strips * n)));
  const int rps = cdiv(H, segs);
  segs = cdiv(H, rps);
  const int per = strips * segs;
  if (int rc = vfi_red_reserve(h, std::max((size_t)65536 * 8, (size_t)n * per * 16))) return rc;
  long long* partial = (long long*)h->red;
  if (fast) laplacian_warp_kernel<<<dim3(cdiv(strips, 4), segs, n), 128, 0, s>>>(rgb, H, W, rps, strips, partial);
  else laplacian_kernel<<<dim3(strips, segs, n), LP_WARPS * 32, 0, s>>>(rgb, H, W, rps, partial);
  VFI_LAUNCHED(h);
  laplacian_finish_kernel<<<n, 256, 0, s>>>(partial, per, (double)H * W, out);
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
#define HM_TW 32     // tile width: the 48-pixel region (tile + 2 x 8 halo) is one 64-bit word per row
#define HM_TH 128    // tile height: tall tiles amortise the per-CTA cell / flag set-up (4x fewer CTAs than 32 x 32)
#define HM_R 8
#define HM_NX (HM_TW + 2 * HM_R)   // 48 region columns (bits)
#define HM_NY (HM_TH + 2 * HM_R)   // 144 region rows (one thread each in the morphology steps)
#define HM_CX 50     // small-mask cells a 48-pixel span can touch when the frame is at least 320 x 180 (ratio >= 1);
#define HM_CY 146    // ... and a 144-pixel span; smaller frames (nearest UPsampling of the thumbnail grid) take the unstaged path
// np.var(last five thumbnails) < thr with numpy's operation order; `newest` >= 0 replaces the ring read of the fifth
// (newest) thumbnail: the fused kernel computes that value itself while another CTA may still be writing it to the ring
__device__ __forceinline__ bool hud_flag_at(const uint8_t* __restrict__ ring, const int* idx, double thr, int i, int newest = -1) {
  int xi[5];
  int S = 0, Q = 0;
#pragma unroll
  for (int k = 0; k < 5; k++) {
    xi[k] = (k == 4 && newest >= 0) ? newest : (int)ring[(size_t)idx[k] * HUD_W * HUD_H + i];
    S += xi[k]; Q += xi[k] * xi[k];
  }
  // 25 * var = 5 * sum(x^2) - (sum x)^2 exactly; numpy's float64 result is within a few ulp of the true variance, so the
  // comparison can only differ from the exact one when the true variance sits (almost) on the threshold: decide those
  // rare cases with numpy's own operation order, everything else with two integer multiplies
  const double v25 = (double)(5 * Q - S * S), t25 = 25.0 * thr;
  if (fabs(v25 - t25) > 1e-6 * (t25 > 1.0 ? t25 : 1.0)) return v25 < t25;
  double x[5];
  double sum = 0.0;
#pragma unroll
  for (int k = 0; k < 5; k++) { x[k] = (double)xi[k]; sum = __dadd_rn(sum, x[k]); }
  const double mean = __ddiv_rn(sum, 5.0);
  double acc = 0.0;
#pragma unroll
  for (int k = 0; k < 5; k++) { const double d = __dsub_rn(x[k], mean); acc = __dadd_rn(acc, __dmul_rn(d, d)); }
  return __ddiv_rn(acc, 5.0) < thr;
}
__device__ __forceinline__ void hm_dilate(unsigned long long* a, unsigned long long* tmp, const bool rowin, unsigned long long inx,
                                          int r, bool complement) {
  const int t = threadIdx.x;
  if (t < HM_NY) {
    unsigned long long w = a[t];
    if (complement) w = ~w;
    w = rowin ? (w & inx) : 0ull;
    unsigned long long hz = w;
    for (int k = 1; k <= r; k++) hz |= (w << k) | (w >> k);
    tmp[t] = hz;
  }
  __syncthreads();
  if (t < HM_NY) {
    unsigned long long v = 0ull;
    for (int k = -r; k <= r; k++)
      if (t + k >= 0 && t + k < HM_NY) v |= tmp[t + k];
    a[t] = complement ? ~v : v;
  }
  __syncthreads();
}
// FUSED (frames of at least 320 x 180): the launch also IS detect_hud's thumbnail step.  Every CTA computes the new frame's
// thumbnail values for the cells it needs straight from the RGB frame (never reading the ring slot being written), the CTA
// whose tile holds a cell's first pixel stores it into the ring, and the last CTA to finish turns the count into the
// coverage and re-arms the two counters: one launch per frame instead of thumbnail + memset + mask + coverage.
// the pixel whose tile's CTA stores cell c of the new thumbnail: every CTA evaluates the same formula, the pixel lies in
// cell c - 1 or c, and a tile's 8-pixel halo always reaches the next cell's first pixel, so exactly one CTA that staged
// cell c owns it
__device__ __forceinline__ int hud_owner_px(int c, double ratio, int n) { return max(0, min(n - 1, (int)floor(__dmul_rn((double)c, ratio)))); }
template <bool FUSED>
__global__ void __launch_bounds__(256) hud_mask_kernel(uint8_t* __restrict__ ring, HudSlots sl, double thr,
                                                       uint8_t* __restrict__ dst, int H, int W, double ify, double ifx,
                                                       unsigned long long* __restrict__ cnt, const uint8_t* __restrict__ rgb,
                                                       double rsy, double rsx, double* __restrict__ coverage_out) {
  __shared__ unsigned long long rows[HM_NY], tmp[HM_NY];
  __shared__ uint8_t flag[HM_CY][HM_CX];
  const int tx0 = blockIdx.x * HM_TW, ty0 = blockIdx.y * HM_TH;
  const int ox = tx0 - HM_R, oy = ty0 - HM_R;
  const int x_lo = max(ox, 0), x_hi = min(ox + HM_NX - 1, W - 1), y_lo = max(oy, 0), y_hi = min(oy + HM_NY - 1, H - 1);
  const int cx0 = min((int)floor(__dmul_rn((double)x_lo, ifx)), HUD_W - 1), cx1 = min((int)floor(__dmul_rn((double)x_hi, ifx)), HUD_W - 1);
  const int cy0 = min((int)floor(__dmul_rn((double)y_lo, ify)), HUD_H - 1), cy1 = min((int)floor(__dmul_rn((double)y_hi, ify)), HUD_H - 1);
  const int ncx = cx1 - cx0 + 1, ncy = cy1 - cy0 + 1;
  const bool staged = ncx <= HM_CX && ncy <= HM_CY;
  if (staged)
    for (int e = threadIdx.x; e < ncx * ncy; e += blockDim.x) {
      const int cy = e / ncx, cx = e - cy * ncx, ax = cx0 + cx, ay = cy0 + cy;
      int newest = -1;
      if (FUSED) {
        newest = resized_gray(rgb, W, lin_tap(ax, rsx, W, true), lin_tap(ay, rsy, H, false));
        const int fx = hud_owner_px(ax, rsx, W), fy = hud_owner_px(ay, rsy, H);
        if (fx >= tx0 && fx < tx0 + HM_TW && fy >= ty0 && fy < ty0 + HM_TH)
          ring[(size_t)sl.idx[4] * HUD_W * HUD_H + ay * HUD_W + ax] = (uint8_t)newest;
      }
      flag[cy][cx] = hud_flag_at(ring, sl.idx, thr, ay * HUD_W + ax, newest) ? 1 : 0;
    }
  // region column / row -> staged cell index (or -1 outside the image): the double multiply + floor once per column / row
  __shared__ short colcell[HM_NX], rowcell[HM_NY];
  if (threadIdx.x < HM_NX) {
    const int gx = ox + (int)threadIdx.x;
    colcell[threadIdx.x] = (gx >= 0 && gx < W) ? (short)(min((int)floor(__dmul_rn((double)gx, ifx)), HUD_W - 1) - cx0) : (short)-1;
  }
  if (threadIdx.x < HM_NY) {
    const int gy = oy + (int)threadIdx.x;
    rowcell[threadIdx.x] = (gy >= 0 && gy < H) ? (short)(min((int)floor(__dmul_rn((double)gy, ify)), HUD_H - 1) - cy0) : (short)-1;
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int t = warp; t < HM_NY; t += 8) {
    const int cy = rowcell[t];
    unsigned int half[2];
#pragma unroll
    for (int hh = 0; hh < 2; hh++) {
      const int c = lane + 32 * hh;
      bool bit = false;
      if (cy >= 0 && c < HM_NX) {
        const int cc = colcell[c];
        if (cc >= 0) bit = staged ? flag[cy][cc] != 0 : hud_flag_at(ring, sl.idx, thr, (cy + cy0) * HUD_W + cc + cx0);
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
  const bool rowin = threadIdx.x < HM_NY && gy_t >= 0 && gy_t < H;
  hm_dilate(rows, tmp, rowin, inx, 2, false);  // MORPH_CLOSE: dilate 5x5 ...
  hm_dilate(rows, tmp, rowin, inx, 4, true);   // ... erode 5x5, then MORPH_OPEN's erode 5x5: together erode 9x9
  hm_dilate(rows, tmp, rowin, inx, 2, false);  // ... dilate 5x5
  // store the tile (4 pixels per thread and step, 32 rows per step) and count it
  unsigned int c4 = 0;
  const int xq = (threadIdx.x & 7) * 4, x = tx0 + xq;
#pragma unroll
  for (int step = 0; step < HM_TH / 32; step++) {
    const int yl = (threadIdx.x >> 3) + 32 * step, y = ty0 + yl;
    if (y < H && x < W) {
      const unsigned long long w = rows[HM_R + yl] >> (HM_R + xq);
      uint8_t v[4];
#pragma unroll
      for (int k = 0; k < 4; k++) { v[k] = (x + k < W) ? (uint8_t)((w >> k) & 1ull) : 0; c4 += v[k]; }
      uint8_t* o = dst + (size_t)y * W + x;
      if (x + 3 < W && ((W & 3) == 0)) *reinterpret_cast<uchar4*>(o) = make_uchar4(v[0], v[1], v[2], v[3]);
      else for (int k = 0; k < 4 && x + k < W; k++) o[k] = v[k];
    }
  }
  __shared__ unsigned long long sh[32];
  const unsigned long long c = block_sum((unsigned long long)c4, sh);
  if (threadIdx.x == 0) {
    if (c) atomicAdd(cnt, c);
    if (FUSED) {
      __threadfence();
      const unsigned long long ticket = atomicAdd(cnt + 1, 1ull);
      if (ticket == (unsigned long long)gridDim.x * gridDim.y - 1) {   // last CTA: every count is in
        __threadfence();
        *coverage_out = (double)atomicAdd(cnt, 0ull) / ((double)H * (double)W);
        cnt[0] = 0; cnt[1] = 0;                                          // re-armed for the next frame
      }
    }
  }
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
  ProfScope ps(h, PC_REDUCE, s);
  if (int rc = vfi_red_reserve(h, std::max((size_t)65536 * 8, npix + HUD_W * HUD_H + 4096))) return rc;
  const long long c = *count;
  // cv2.resize(gray, (320, 180)): inv_scale = dsize / ssize, scale = 1 / inv_scale (doubles, as OpenCV computes them)
  const double rsx = 1.0 / ((double)HUD_W / (double)W), rsy = 1.0 / ((double)HUD_H / (double)H);
  HudSlots sl;
  for (int k = 0; k < 5; k++) sl.idx[k] = (int)(((c - 4 + k) % hist_len + hist_len) % hist_len);  // oldest -> newest of the last five
  const double ifx = 1.0 / ((double)W / (double)HUD_W), ify = 1.0 / ((double)H / (double)HUD_H);
  if (c + 1 >= 5 && W >= HUD_W && H >= HUD_H) {  // steady state: one launch (thumbnail + mask + coverage)
    hud_mask_kernel<true><<<dim3(cdiv(W, HM_TW), cdiv(H, HM_TH)), 256, 0, s>>>(hist_ring, sl, var_threshold, mask_out, H, W, ify, ifx,
                                                                             h->hud_cnt, rgb, rsy, rsx, coverage_out);
    VFI_LAUNCHED(h);
    *count = c + 1;
    return VFI_OK;
  }
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
  VFI_CUDA(h, cudaMemsetAsync(cnt, 0, 8, s));
  hud_mask_kernel<false><<<dim3(cdiv(W, HM_TW), cdiv(H, HM_TH)), 256, 0, s>>>(hist_ring, sl, var_threshold, mask_out, H, W, ify, ifx, cnt,
                                                                            nullptr, 0.0, 0.0, nullptr);
  VFI_LAUNCHED(h);
  coverage_finish_kernel<<<1, 1, 0, s>>>(cnt, (double)npix, coverage_out);
  VFI_LAUNCHED(h);
  return VFI_OK;
}

// ---------------------------------------------------------------------------------------------
// The experiment runner's motion-threshold router (scripts/run_experiments.py:848-874): gray thumbnails and the linear
// blend it falls back to.
//   resize_gray : cv2.resize(cv2.cvtColor(frame, *2GRAY), (dw, dh)) -- INTER_LINEAR, uint8 fixed point (lin_tap above)
//   add_weighted: cv2.addWeighted(a, alpha, b, beta, 0) for uint8: saturate_cast<uchar>(cvRound(a*alpha + b*beta)) in float
template <int CH, bool BGR>
__device__ __forceinline__ int gray_src(const uint8_t* __restrict__ img, int W, int y, int x) {
  if (CH == 1) return img[(size_t)y * W + x];
  const uint8_t* q = img + ((size_t)y * W + x) * 3;
  return BGR ? gray_u8(q[2], q[1], q[0]) : gray_u8(q[0], q[1], q[2]);
}
template <int CH, bool BGR>
__global__ void resize_gray_kernel(const uint8_t* __restrict__ src, int H, int W, int dh, int dw, double sy, double sx,
                                   uint8_t* __restrict__ dst) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= dw) return;
  const uint8_t* img = src + (size_t)blockIdx.z * H * W * CH;
  int v;
  if (dh == H && dw == W) v = gray_src<CH, BGR>(img, W, y, x);
  else {
    const LinTap tx = lin_tap(x, sx, W, true), ty = lin_tap(y, sy, H, false);
    const int h0 = gray_src<CH, BGR>(img, W, ty.s0, tx.s0) * tx.a0 + gray_src<CH, BGR>(img, W, ty.s0, tx.s1) * tx.a1;
    const int h1 = gray_src<CH, BGR>(img, W, ty.s1, tx.s0) * tx.a0 + gray_src<CH, BGR>(img, W, ty.s1, tx.s1) * tx.a1;
    v = (((ty.a0 * (h0 >> 4)) >> 16) + ((ty.a1 * (h1 >> 4)) >> 16) + 2) >> 2;
  }
  dst[((size_t)blockIdx.z * dh + y) * dw + x] = (uint8_t)v;
}
extern "C" __attribute__((visibility("default"))) int vfi_resize_gray_u8(vfi_handle* h, const uint8_t* src, int n, int H, int W, int channels,
                                                                           int bgr, int dh, int dw, uint8_t* dst, void* stream) {
  if (!h) return VFI_E_ARG;
  if (!src || !dst || n < 1 || n > 65535 || H < 1 || W < 1 || dh < 1 || dw < 1 || dh > 65535 || (channels != 1 && channels != 3))
    VFI_FAIL(h, VFI_E_ARG, "bad argument");
  VFI_ENTER(h, stream);
  cudaStream_t s = (cudaStream_t)stream;
  const double sx = 1.0 / ((double)dw / (double)W), sy = 1.0 / ((double)dh / (double)H);   // as cv::resize computes them
  const dim3 grid(cdiv(dw, 128), dh, n);
  ProfScope ps(h, PC_REDUCE, s);
  if (channels == 1) resize_gray_kernel<1, false><<<grid, 128, 0, s>>>(src, H, W, dh, dw, sy, sx, dst);
  else if (bgr) resize_gray_kernel<3, true><<<grid, 128, 0, s>>>(src, H, W, dh, dw, sy, sx, dst);
  else resize_gray_kernel<3, false><<<grid, 128, 0, s>>>(src, H, W, dh, dw, sy, sx, dst);
  VFI_LAUNCHED(h);
  return VFI_OK;
}
__global__ void add_weighted_kernel(const uint8_t* __restrict__ a, const uint8_t* __restrict__ b, size_t elems, float alpha, float beta,
                                    uint8_t* __restrict__ out) {
  const bool aligned = ((((uintptr_t)a) | ((uintptr_t)b) | ((uintptr_t)out)) & 15) == 0;
  const size_t nvec = aligned ? elems / 16 : 0;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < nvec; i += (size_t)gridDim.x * blockDim.x) {
    const uint4 x = __ldg(reinterpret_cast<const uint4*>(a) + i), y = __ldg(reinterpret_cast<const uint4*>(b) + i);
    const uint32_t xs[4] = {x.x, x.y, x.z, x.w}, ys[4] = {y.x, y.y, y.z, y.w};
    uint32_t o[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
      uint32_t w = 0;
#pragma unroll
      for (int j = 0; j < 4; j++) {
        const float v = fmaf((float)((xs[k] >> (8 * j)) & 0xff), alpha, __fmul_rn((float)((ys[k] >> (8 * j)) & 0xff), beta));
        w |= (uint32_t)min(max(__float2int_rn(v), 0), 255) << (8 * j);
      }
      o[k] = w;
    }
    reinterpret_cast<uint4*>(out)[i] = make_uint4(o[0], o[1], o[2], o[3]);
  }
  for (size_t i = nvec * 16 + blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < elems; i += (size_t)gridDim.x * blockDim.x) {
    const float v = fmaf((float)a[i], alpha, __fmul_rn((float)b[i], beta));
    out[i] = (uint8_t)min(max(__float2int_rn(v), 0), 255);
  }
}
extern "C" __attribute__((visibility("default"))) int vfi_add_weighted_u8(vfi_handle* h, const uint8_t* a, const uint8_t* b, size_t elems,
                                                                            double alpha, double beta, uint8_t* out, void* stream) {
  if (!h) return VFI_E_ARG;
  if (!a || !b || !out || elems < 1) VFI_FAIL(h, VFI_E_ARG, "bad argument");
  VFI_ENTER(h, stream);
  cudaStream_t s = (cudaStream_t)stream;
  ProfScope ps(h, PC_BLEND, s);
  const int blocks = (int)std::min<size_t>((size_t)h->sm_count * 16, (elems / 16 + 255) / 256 + 1);
  add_weighted_kernel<<<blocks, 256, 0, s>>>(a, b, elems, (float)alpha, (float)beta, out);
  VFI_LAUNCHED(h);
  return VFI_OK;
}

// ---------------------------------------------------------------------------------------------
// cv2.Canny(gray, low, high) (aperture 3, L1 gradient) -- the edge-density term of the MCAR complexity score,
// scripts/run_experiments.py:466-468.  OpenCV's algorithm (imgproc/canny.cpp), all integer, so the edge map is bit-exact:
//   Sobel 3x3 dx, dy (BORDER_REPLICATE), mag = |dx| + |dy| (0 outside the image);
//   non-maximum suppression along the gradient direction quantised with tan(22.5 deg) in 15-bit fixed point;
//   label 2 (strong: local maximum, mag > high), 0 (weak: local maximum, low < mag <= high), 1 (no edge);
//   hysteresis: weak pixels 8-connected to a strong one become strong -- tile-local fixed point in shared memory,
//   relaunched until no tile changes (the result does not depend on the order).
#define CN_T 32
__global__ void __launch_bounds__(256) canny_nms_kernel(const uint8_t* __restrict__ gray, int H, int W, int low, int high,
                                                        uint8_t* __restrict__ label) {
  __shared__ uint8_t g[CN_T + 4][CN_T + 4];
  __shared__ short sdx[CN_T + 2][CN_T + 2], sdy[CN_T + 2][CN_T + 2];
  __shared__ int mg[CN_T + 2][CN_T + 2];
  const uint8_t* img = gray + (size_t)blockIdx.z * H * W;
  const int x0 = blockIdx.x * CN_T, y0 = blockIdx.y * CN_T;
  for (int e = threadIdx.x; e < (CN_T + 4) * (CN_T + 4); e += blockDim.x) {
    const int ly = e / (CN_T + 4), lx = e - ly * (CN_T + 4);
    const int gy = min(max(y0 + ly - 2, 0), H - 1), gx = min(max(x0 + lx - 2, 0), W - 1);   // BORDER_REPLICATE
    g[ly][lx] = img[(size_t)gy * W + gx];
  }
  __syncthreads();
  for (int e = threadIdx.x; e < (CN_T + 2) * (CN_T + 2); e += blockDim.x) {
    const int ly = e / (CN_T + 2), lx = e - ly * (CN_T + 2);
    const int gy = y0 + ly - 1, gx = x0 + lx - 1;
    int dx = 0, dy = 0, m = 0;
    if (gy >= 0 && gy < H && gx >= 0 && gx < W) {
      const int a = g[ly][lx], b = g[ly][lx + 1], c = g[ly][lx + 2];
      const int d = g[ly + 1][lx], f = g[ly + 1][lx + 2];
      const int p = g[ly + 2][lx], q = g[ly + 2][lx + 1], r = g[ly + 2][lx + 2];
      dx = (c - a) + 2 * (f - d) + (r - p);
      dy = (p - a) + 2 * (q - b) + (r - c);
      m = abs(dx) + abs(dy);
    }
    sdx[ly][lx] = (short)dx; sdy[ly][lx] = (short)dy; mg[ly][lx] = m;
  }
  __syncthreads();
  for (int e = threadIdx.x; e < CN_T * CN_T; e += blockDim.x) {
    const int ly = (e >> 5) + 1, lx = (e & 31) + 1;
    const int gy = y0 + ly - 1, gx = x0 + lx - 1;
    if (gy >= H || gx >= W) continue;
    const int m = mg[ly][lx];
    uint8_t lab = 1;
    if (m > low) {
      const int xs = sdx[ly][lx], ys = sdy[ly][lx];
      const int ax = abs(xs), ay = abs(ys) << 15;
      const int tg22x = ax * 13573;
      bool ismax;
      if (ay < tg22x) ismax = m > mg[ly][lx - 1] && m >= mg[ly][lx + 1];
      else {
        const int tg67x = tg22x + (ax << 16);
        if (ay > tg67x) ismax = m > mg[ly - 1][lx] && m >= mg[ly + 1][lx];
        else {
          const int sg = (xs ^ ys) < 0 ? -1 : 1;
          ismax = m > mg[ly - 1][lx - sg] && m > mg[ly + 1][lx + sg];
        }
      }
      if (ismax) lab = m > high ? 2 : 0;
    }
    label[((size_t)blockIdx.z * H + gy) * W + gx] = lab;
  }
}
__global__ void __launch_bounds__(256) canny_hyst_kernel(uint8_t* __restrict__ label, int H, int W, int* __restrict__ changed) {
  __shared__ uint8_t t[CN_T + 2][CN_T + 2];
  __shared__ int again, any;
  uint8_t* lab = label + (size_t)blockIdx.z * H * W;
  const int x0 = blockIdx.x * CN_T, y0 = blockIdx.y * CN_T;
  for (int e = threadIdx.x; e < (CN_T + 2) * (CN_T + 2); e += blockDim.x) {
    const int ly = e / (CN_T + 2), lx = e - ly * (CN_T + 2);
    const int gy = y0 + ly - 1, gx = x0 + lx - 1;
    t[ly][lx] = (gy >= 0 && gy < H && gx >= 0 && gx < W) ? lab[(size_t)gy * W + gx] : 1;
  }
  if (threadIdx.x == 0) any = 0;
  __syncthreads();
  for (;;) {
    if (threadIdx.x == 0) again = 0;
    __syncthreads();
    for (int e = threadIdx.x; e < CN_T * CN_T; e += blockDim.x) {
      const int ly = (e >> 5) + 1, lx = (e & 31) + 1;
      if (t[ly][lx] == 0) {
        const bool s = t[ly - 1][lx - 1] == 2 || t[ly - 1][lx] == 2 || t[ly - 1][lx + 1] == 2 || t[ly][lx - 1] == 2 ||
                       t[ly][lx + 1] == 2 || t[ly + 1][lx - 1] == 2 || t[ly + 1][lx] == 2 || t[ly + 1][lx + 1] == 2;
        if (s) { t[ly][lx] = 2; again = 1; }       // monotone 0 -> 2: racing readers only see it earlier or later
      }
    }
    __syncthreads();
    if (!again) break;
    if (threadIdx.x == 0) any = 1;
    __syncthreads();
  }
  if (any) {
    for (int e = threadIdx.x; e < CN_T * CN_T; e += blockDim.x) {
      const int ly = (e >> 5) + 1, lx = (e & 31) + 1;
      const int gy = y0 + ly - 1, gx = x0 + lx - 1;
      if (gy < H && gx < W && t[ly][lx] == 2) lab[(size_t)gy * W + gx] = 2;
    }
    if (threadIdx.x == 0) *changed = 1;
  }
}
__global__ void canny_finish_kernel(const uint8_t* __restrict__ label, size_t px, uint8_t* __restrict__ edges, unsigned long long* __restrict__ cnt) {
  __shared__ unsigned long long sh[32];
  const int f = blockIdx.y;
  unsigned long long c = 0;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < px; i += (size_t)gridDim.x * blockDim.x) {
    const bool e = label[(size_t)f * px + i] == 2;
    if (edges) edges[(size_t)f * px + i] = e ? 255 : 0;
    c += e;
  }
  c = block_sum(c, sh);
  if (threadIdx.x == 0 && c) atomicAdd(cnt + f, c);
}
__global__ void canny_density_kernel(const unsigned long long* cnt, int n, double px, double* out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = (double)cnt[i] / px;
}
extern "C" __attribute__((visibility("default"))) int vfi_canny_u8(vfi_handle* h, const uint8_t* gray, int n, int H, int W, double low, double high,
                                                                     uint8_t* edges_out, double* density_out, void* stream) {
  if (!h) return VFI_E_ARG;
  if (!gray || (!edges_out && !density_out) || n < 1 || n > 4096 || H < 1 || W < 1) VFI_FAIL(h, VFI_E_ARG, "bad argument");
  VFI_ENTER(h, stream);
  cudaStream_t s = (cudaStream_t)stream;
  if (low > high) std::swap(low, high);
  const int ilow = (int)floor(low), ihigh = (int)floor(high);
  const size_t px = (size_t)H * W;
  if (int rc = vfi_red_reserve(h, std::max((size_t)65536 * 8, (size_t)n * px + (size_t)n * 8 + 256))) return rc;
  uint8_t* label = (uint8_t*)h->red;
  unsigned long long* cnt = (unsigned long long*)(label + (((size_t)n * px + 63) & ~(size_t)63));
  int* changed = (int*)(cnt + n);
  const dim3 grid(cdiv(W, CN_T), cdiv(H, CN_T), n);
  ProfScope ps(h, PC_REDUCE, s);
  canny_nms_kernel<<<grid, 256, 0, s>>>(gray, H, W, ilow, ihigh, label);
  VFI_LAUNCHED(h);
  for (int it = 0; it < 4096; it++) {          // each pass moves the frontier at least one tile; stops at the fixed point
    VFI_CUDA(h, cudaMemsetAsync(changed, 0, sizeof(int), s));
    canny_hyst_kernel<<<grid, 256, 0, s>>>(label, H, W, changed);
    VFI_LAUNCHED(h);
    int host = 0;
    VFI_CUDA(h, cudaMemcpyAsync(&host, changed, sizeof(int), cudaMemcpyDeviceToHost, s));
    VFI_CUDA(h, cudaStreamSynchronize(s));
    if (!host) break;
  }
  VFI_CUDA(h, cudaMemsetAsync(cnt, 0, (size_t)n * 8, s));
  canny_finish_kernel<<<dim3(std::min<size_t>(296, cdiv(px, 1024)), n), 256, 0, s>>>(label, px, edges_out, cnt);
  VFI_LAUNCHED(h);
  if (density_out) {
    canny_density_kernel<<<cdiv(n, 128), 128, 0, s>>>(cnt, n, (double)px, density_out);
    VFI_LAUNCHED(h);
  }
  return VFI_OK;
}

// ---------------------------------------------------------------------------------------------
// MCAR tier 2 (scripts/run_experiments.py:498-503, :526-531): cv2.remap(prev, x + flow_x * t, y + flow_y * t, INTER_LINEAR)
// on uint8 frames (BORDER_CONSTANT 0).  OpenCV's fixed-point path restated exactly: the float32 map coordinate is rounded to
// 1/32 pixel (cvRound(v * 32): half to even), the four bilinear weights are the 5-bit fractions' products scaled to 2^15
// ((32-a)(32-b), a(32-b), (32-a)b, ab times 32: they sum to 2^15 with no correction), result = (sum + 2^14) >> 15.
__global__ void remap_flow_kernel(const uint8_t* __restrict__ src, const float2* __restrict__ flow, float t, int H, int W, int C,
                                  uint8_t* __restrict__ dst) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= W) return;
  const float2 f = __ldg(flow + (size_t)y * W + x);
  const float mx = __fadd_rn((float)x, __fmul_rn(f.x, t)), my = __fadd_rn((float)y, __fmul_rn(f.y, t));
  const int sx = __float2int_rn(__fmul_rn(mx, 32.f)), sy = __float2int_rn(__fmul_rn(my, 32.f));
  const int ix = sx >> 5, iy = sy >> 5, a = sx & 31, b = sy & 31;
  const int w00 = (32 - a) * (32 - b), w01 = a * (32 - b), w10 = (32 - a) * b, w11 = a * b;
  const bool x0 = (unsigned)ix < (unsigned)W, x1 = (unsigned)(ix + 1) < (unsigned)W;
  const bool y0 = (unsigned)iy < (unsigned)H, y1 = (unsigned)(iy + 1) < (unsigned)H;
  const uint8_t* p = src + ((ptrdiff_t)iy * W + ix) * C;
  uint8_t* o = dst + ((size_t)y * W + x) * C;
  for (int c = 0; c < C; c++) {
    const int p00 = (x0 && y0) ? p[c] : 0, p01 = (x1 && y0) ? p[C + c] : 0;
    const int p10 = (x0 && y1) ? p[(size_t)W * C + c] : 0, p11 = (x1 && y1) ? p[(size_t)W * C + C + c] : 0;
    o[c] = (uint8_t)(((p00 * w00 + p01 * w01 + p10 * w10 + p11 * w11) * 32 + 16384) >> 15);
  }
}
extern "C" __attribute__((visibility("default"))) int vfi_remap_flow_u8(vfi_handle* h, const uint8_t* src, const float* flow, double t, int H, int W,
                                                                          int channels, uint8_t* dst, void* stream) {
  if (!h) return VFI_E_ARG;
  if (!src || !flow || !dst || H < 1 || W < 1 || H > 65535 || channels < 1 || channels > 4) VFI_FAIL(h, VFI_E_ARG, "bad argument");
  VFI_ENTER(h, stream);
  cudaStream_t s = (cudaStream_t)stream;
  ProfScope ps(h, PC_BLEND, s);
  remap_flow_kernel<<<dim3(cdiv(W, 128), H), 128, 0, s>>>(src, reinterpret_cast<const float2*>(flow), (float)t, H, W, channels, dst);
  VFI_LAUNCHED(h);
  return VFI_OK;
}

// ---------------------------------------------------------------------------------------------
// Planar YUV 4:2:0 (I420) result format: cv2.cvtColor(frame, COLOR_RGB2YUV_I420 / COLOR_BGR2YUV_I420) restated exactly
// (OpenCV imgproc color_yuv: ITU-R BT.601 limited range, 20-bit fixed point, chroma taken from the TOP-LEFT pixel of each
// 2 x 2 block, no averaging; pinned against cv2 4.13 by tests/test_gpu_reduce.py).  1.5 bytes per pixel instead of 3: the
// format the experiment runner's encoder wants (scripts/run_experiments.py pipes raw frames into libx264, which converts to
// yuv420p), and half the device -> host bytes of the end-to-end path, which is host-link bound at 4 and 8 GPUs per box.
// Layout per frame, as cv2 returns it: (H * 3 / 2) x W bytes = Y plane H x W, U plane (H/2) x (W/2), V plane (H/2) x (W/2).
#define YUV_SHIFT 20
__device__ __forceinline__ int yuv_y(int r, int g, int b) { return (269484 * r + 528482 * g + 102760 * b + (16 << YUV_SHIFT) + (1 << (YUV_SHIFT - 1))) >> YUV_SHIFT; }
__device__ __forceinline__ int yuv_u(int r, int g, int b) { return (-155188 * r - 305135 * g + 460324 * b + (128 << YUV_SHIFT) + (1 << (YUV_SHIFT - 1))) >> YUV_SHIFT; }
__device__ __forceinline__ int yuv_v(int r, int g, int b) { return (460324 * r - 385875 * g - 74448 * b + (128 << YUV_SHIFT) + (1 << (YUV_SHIFT - 1))) >> YUV_SHIFT; }
// FAST (W % 4 == 0, 4-byte aligned frames): a thread converts 4 x 2 pixels from six aligned 32-bit words into two 32-bit
// luma stores and one 16-bit store per chroma plane; otherwise a thread converts one 2 x 2 block with byte accesses.
template <bool FAST, bool BGR>
__global__ void rgb_to_i420_kernel(const uint8_t* __restrict__ src, int H, int W, uint8_t* __restrict__ dst) {
  const int f = blockIdx.z, y = (blockIdx.y * blockDim.y + threadIdx.y) * 2;
  const int x = (blockIdx.x * blockDim.x + threadIdx.x) * (FAST ? 4 : 2);
  if (x >= W || y >= H) return;
  const uint8_t* img = src + (size_t)f * H * W * 3;
  uint8_t* Y = dst + (size_t)f * H * W * 3 / 2;
  uint8_t* U = Y + (size_t)H * W;
  uint8_t* V = U + (size_t)(H / 2) * (W / 2);
  constexpr int NP = FAST ? 4 : 2;
  int uu[NP / 2], vv[NP / 2];
#pragma unroll
  for (int j = 0; j < 2; j++) {
    uint32_t px[NP];
    const uint8_t* row = img + ((size_t)(y + j) * W + x) * 3;
    if (FAST) {
      const uint32_t* p = reinterpret_cast<const uint32_t*>(row);
      const uint32_t w0 = __ldg(p), w1 = __ldg(p + 1), w2 = __ldg(p + 2);
      px[0] = w0; px[1] = __byte_perm(w0, w1, 0x0543); px[2] = __byte_perm(w1, w2, 0x0432); px[NP - 1] = w2 >> 8;
    } else {
#pragma unroll
      for (int i = 0; i < NP; i++) px[i] = (uint32_t)row[3 * i] | ((uint32_t)row[3 * i + 1] << 8) | ((uint32_t)row[3 * i + 2] << 16);
    }
    uint32_t yw = 0;
#pragma unroll
    for (int i = 0; i < NP; i++) {
      const int c0 = px[i] & 0xff, c1 = (px[i] >> 8) & 0xff, c2 = (px[i] >> 16) & 0xff;
      const int r = BGR ? c2 : c0, b = BGR ? c0 : c2;
      yw |= (uint32_t)yuv_y(r, c1, b) << (8 * i);
      if (j == 0 && (i & 1) == 0) { uu[i / 2] = yuv_u(r, c1, b); vv[i / 2] = yuv_v(r, c1, b); }
    }
    uint8_t* yo = Y + (size_t)(y + j) * W + x;
    if (FAST) *reinterpret_cast<uint32_t*>(yo) = yw;
    else { yo[0] = (uint8_t)(yw & 0xff); yo[1] = (uint8_t)(yw >> 8); }
  }
  const size_t co = (size_t)(y / 2) * (W / 2) + x / 2;
  if (FAST) {
    *reinterpret_cast<uint16_t*>(U + co) = (uint16_t)(uu[0] | (uu[NP / 2 - 1] << 8));
    *reinterpret_cast<uint16_t*>(V + co) = (uint16_t)(vv[0] | (vv[NP / 2 - 1] << 8));
  } else { U[co] = (uint8_t)uu[0]; V[co] = (uint8_t)vv[0]; }
}
extern "C" __attribute__((visibility("default"))) int vfi_rgb_to_yuv420_u8(vfi_handle* h, const uint8_t* src, int n, int H, int W, int bgr,
                                                                             uint8_t* dst, void* stream) {
  if (!h) return VFI_E_ARG;
  if (!src || !dst || n < 1 || n > 65535 || H < 2 || W < 2) VFI_FAIL(h, VFI_E_ARG, "bad argument");
  if ((H & 1) || (W & 1)) VFI_FAIL(h, VFI_E_ARG, "4:2:0 needs even frame sizes, got %dx%d", H, W);
  VFI_ENTER(h, stream);
  cudaStream_t s = (cudaStream_t)stream;
  ProfScope ps(h, PC_MISC, s);
  // per-frame plane offsets stay 4-byte aligned when H * W * 3 / 2 is a multiple of 4: W % 4 == 0 and H even
  const bool fast = (W & 3) == 0 && ((((uintptr_t)src) | ((uintptr_t)dst)) & 3) == 0;
  const dim3 block(32, 4), grid(cdiv(W / (fast ? 4 : 2), 32), cdiv(H / 2, 4), n);
  if (fast) { if (bgr) rgb_to_i420_kernel<true, true><<<grid, block, 0, s>>>(src, H, W, dst); else rgb_to_i420_kernel<true, false><<<grid, block, 0, s>>>(src, H, W, dst); }
  else { if (bgr) rgb_to_i420_kernel<false, true><<<grid, block, 0, s>>>(src, H, W, dst); else rgb_to_i420_kernel<false, false><<<grid, block, 0, s>>>(src, H, W, dst); }
  VFI_LAUNCHED(h);
  return VFI_OK;
}

// ---------------------------------------------------------------------------------------------
// vfi_psnr_u8 without any allocation (no counterpart in the reference).
// vfi_psnr_u8 keeps its per-pair SSE accumulators in the handle's cudaMalloc'd scratch (vfi_red_reserve); callers that may
// not allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc) use this variant instead: the
// accumulators live in a file-scope __device__ array of VFI_NOALLOC_MAX_PAIRS entries (32 KB, held by every GPU that loads the
// library until exit).  In HOST code `g_noalloc_sse` is only a stand-in for the symbol, not a device address: the address is
// resolved once per device, with that device current, by cudaGetSymbolAddress and cached.  Every call shares the array and
// finds the previous call's sums in it: the entries it uses are cleared first (cudaMemsetAsync on `stream`), and calls that
// overlap on the GPU (different streams, no ordering) race on it.
static __device__ __align__(256) unsigned long long g_noalloc_sse[VFI_NOALLOC_MAX_PAIRS];
static unsigned long long* g_noalloc_ptr[64] = {nullptr};   // per device: the array's DEVICE address

extern "C" __attribute__((visibility("default"))) int vfi_psnr_u8_noalloc(vfi_handle* h, const uint8_t* a, const uint8_t* b, int n, size_t elems,
                                                                            unsigned long long* sse_out, double* psnr_out, void* stream) {
  if (!h) return VFI_E_ARG;
  if (!a || !b || n < 1 || elems < 1 || (!sse_out && !psnr_out)) VFI_FAIL(h, VFI_E_ARG, "bad argument");
  if (n > VFI_NOALLOC_MAX_PAIRS)
    VFI_FAIL(h, VFI_E_ARG, "vfi_psnr_u8_noalloc: %d frame pairs exceed the static scratch of %d; split the call or use vfi_psnr_u8", n,
             VFI_NOALLOC_MAX_PAIRS);
  if (h->device < 0 || h->device >= 64) VFI_FAIL(h, VFI_E_ARG, "device index %d out of range", h->device);
  // the handle's device current for the call, the caller's restored afterwards; the handle's own scratch is not touched, so
  // none of its cross-stream ordering is needed
  int prev = -1;
  if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
  if (prev != h->device && cudaSetDevice(h->device) != cudaSuccess) { cudaGetLastError(); VFI_FAIL(h, VFI_E_CUDA, "cudaSetDevice(%d) failed", h->device); }
  int rc = VFI_OK;
  cudaStream_t s = (cudaStream_t)stream;
  unsigned long long* acc = g_noalloc_ptr[h->device];
  if (!acc) {   // set-up, once per device (do the first call outside a graph capture)
    const cudaError_t e = cudaGetSymbolAddress((void**)&acc, g_noalloc_sse);
    if (e != cudaSuccess || !acc) {
      cudaGetLastError();
      char msg[256];
      snprintf(msg, sizeof(msg), "cudaGetSymbolAddress(g_noalloc_sse) failed: %s", cudaGetErrorString(e));
      h->err = msg;
      rc = VFI_E_CUDA;
    } else {
      g_noalloc_ptr[h->device] = acc;
    }
  }
  if (rc == VFI_OK) {
    cudaError_t e = cudaMemsetAsync(acc, 0, (size_t)n * sizeof(unsigned long long), s);   // the last call's sums are still there
    if (e == cudaSuccess) {
      int bx = (int)std::min<size_t>(PSNR_BLOCKS_PER_FRAME, (elems / 16 + 255) / 256 + 1);
      if (n >= 64) bx = std::min(bx, 32);
      psnr_sse_kernel<<<dim3(bx, n), 256, 0, s>>>(a, b, elems, acc);
      h->launches++;
      psnr_finish_kernel<<<cdiv(n, 128), 128, 0, s>>>(acc, elems, n, sse_out, psnr_out);
      h->launches++;
      e = cudaGetLastError();
    }
    if (e != cudaSuccess) { h->err = std::string("vfi_psnr_u8_noalloc: ") + cudaGetErrorString(e); rc = VFI_E_CUDA; }
  }
  if (prev >= 0 && prev != h->device) cudaSetDevice(prev);
  return rc;
}
