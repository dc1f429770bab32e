// farneback.cu -- dense Farneback optical flow on the device.
//
// Replaces cv2.calcOpticalFlowFarneback(g0, g1, None, 0.5, 3, 15, 3, 5, 1.2, 0) at
// models/novel/adaptive_pipeline.py:119-128 (and the (0.5, 1, 15, 1, 5, 1.1, 0) variant of
// scripts/run_experiments.py:359-362).  The algorithm is OpenCV's (modules/video/src/optflowgf.cpp; third-party, restated
// and pinned against cv2 in oracle/farneback_ref.py), laid out for the GPU:
//   per pyramid level, coarsest first:
//     fb_blur_h / fb_blur_v   GaussianBlur(full-res gray) + resize(INTER_LINEAR) to the level, separable, fused per axis
//     fb_polyexp              FarnebackPolyExp: 2n+1 vertical taps (fp32) then horizontal taps (fp64), smem tiles
//     fb_matrices             FarnebackUpdateMatrices from the previous level's flow (upsampled x 1/pyr_scale on the fly)
//     fb_blur_solve x iters   winsize x winsize clamped box sums of the 5 matrix planes in fp64 (vertical sliding sums in
//                             a shared-memory tile, horizontal sliding sums in registers), the 2x2 solve, and -- except on
//                             the last iteration -- the next UpdateMatrices fused behind it (M ping-pong)
// All planes are planar fp32 [image][channel][h][w]; nothing but the final flow (and optional statistics) leaves the GPU.
// This file is compiled with -fmad=false: OpenCV's scalar code rounds every product and sum separately.
#include <math.h>

#include "common.cuh"

namespace {

#define FB_MAX_KR 64    // Gaussian pyramid kernel radius
#define FB_MAX_N 7      // poly_n
#define FB_MAX_M 15     // winsize / 2

struct FbKern { int r; float k[FB_MAX_KR + 1]; };  // k[d] = tap at distance d from the centre (symmetric)
struct FbPoly { int n; float g[FB_MAX_N + 1], xg[FB_MAX_N + 1], xxg[FB_MAX_N + 1]; double ig11, ig03, ig33, ig55; };

__device__ __forceinline__ int refl101(int i, int n) {
  if (i < 0) i = -i;
  if (i >= n) i = 2 * (n - 1) - i;
  return min(max(i, 0), n - 1);
}
// cv::resize INTER_LINEAR source taps of destination index d (imgproc/resize.cpp), float weights
struct FTap { int s0, s1; float a0, a1; };
__device__ __forceinline__ FTap ftap(int d, double scale, int src, bool clamp_frac) {
  float f = (float)(((double)d + 0.5) * scale - 0.5);
  int s = (int)floorf(f);
  f = f - (float)s;
  if (clamp_frac) {
    if (s < 0) { s = 0; f = 0.f; }
    if (s >= src - 1) { s = src - 1; f = 0.f; }
  }
  FTap t;
  t.a0 = 1.f - f; t.a1 = f;
  t.s0 = min(max(s, 0), src - 1);
  t.s1 = min(max(s + 1, 0), src - 1);
  return t;
}

// ---- gray ---------------------------------------------------------------------------------------------------------
__global__ void fb_gray_kernel(const uint8_t* __restrict__ img0, const uint8_t* __restrict__ img1, size_t px, int channels,
                               uint8_t* __restrict__ gray /*[n][2][px]*/) {
  const int f = blockIdx.y & 1, i = blockIdx.y >> 1;
  const uint8_t* src = (f ? img1 : img0) + (size_t)i * px * channels;
  uint8_t* dst = gray + ((size_t)blockIdx.y) * px;
  for (size_t p = blockIdx.x * (size_t)blockDim.x + threadIdx.x; p < px; p += (size_t)gridDim.x * blockDim.x) {
    if (channels == 1) dst[p] = src[p];
    else {
      const uint8_t* q = src + p * 3;
      dst[p] = (uint8_t)((q[0] * 9798 + q[1] * 19235 + q[2] * 3735 + 16384) >> 15);  // cv2 RGB2GRAY
    }
  }
}

// ---- Gaussian pyramid level: horizontal blur + horizontal resize, then vertical blur + vertical resize -------------
__global__ void fb_blur_h_kernel(const uint8_t* __restrict__ gray, float* __restrict__ A, int H, int W, int w, double scale_x,
                                 FbKern kk) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= w) return;
  const uint8_t* g = gray + ((size_t)blockIdx.z * H + y) * W;
  FTap t;
  if (w == W) { t.s0 = t.s1 = x; t.a0 = 1.f; t.a1 = 0.f; } else t = ftap(x, scale_x, W, true);
  float v[2];
#pragma unroll
  for (int j = 0; j < 2; j++) {
    const int s = j ? t.s1 : t.s0;
    if (j && t.a1 == 0.f) { v[1] = 0.f; break; }
    float acc = kk.k[0] * (float)g[s];
    for (int d = 1; d <= kk.r; d++) acc = acc + kk.k[d] * ((float)g[refl101(s - d, W)] + (float)g[refl101(s + d, W)]);
    v[j] = acc;
  }
  A[((size_t)blockIdx.z * H + y) * w + x] = t.a1 == 0.f ? v[0] : v[0] * t.a0 + v[1] * t.a1;
}
__global__ void fb_blur_v_kernel(const float* __restrict__ A, float* __restrict__ I, int H, int h, int w, double scale_y,
                                 FbKern kk) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= w) return;
  const float* a = A + (size_t)blockIdx.z * H * w + x;
  FTap t;
  if (h == H) { t.s0 = t.s1 = y; t.a0 = 1.f; t.a1 = 0.f; } else t = ftap(y, scale_y, H, false);
  float v[2];
#pragma unroll
  for (int j = 0; j < 2; j++) {
    const int s = j ? t.s1 : t.s0;
    if (j && t.a1 == 0.f) { v[1] = 0.f; break; }
    float acc = kk.k[0] * a[(size_t)s * w];
    for (int d = 1; d <= kk.r; d++) acc = acc + kk.k[d] * (a[(size_t)refl101(s - d, H) * w] + a[(size_t)refl101(s + d, H) * w]);
    v[j] = acc;
  }
  I[((size_t)blockIdx.z * h + y) * w + x] = t.a1 == 0.f ? v[0] : v[0] * t.a0 + v[1] * t.a1;
}

// Level 0 (no resize, 3-tap kernel): both passes in one launch, same operation order as fb_blur_h then fb_blur_v
// (horizontal taps first, then vertical), so the plane is bit-identical; each thread reads its 3 x 3 gray bytes (L1 hits).
__global__ void fb_blur3_kernel(const uint8_t* __restrict__ gray, float* __restrict__ I, int H, int W, float k0, float k1) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= W) return;
  const uint8_t* g = gray + (size_t)blockIdx.z * H * W;
  const int xl = refl101(x - 1, W), xr = refl101(x + 1, W);
  float a[3];
#pragma unroll
  for (int j = 0; j < 3; j++) {
    const uint8_t* row = g + (size_t)refl101(y - 1 + j, H) * W;
    a[j] = k0 * (float)row[x] + k1 * ((float)row[xl] + (float)row[xr]);
  }
  I[((size_t)blockIdx.z * H + y) * W + x] = k0 * a[1] + k1 * (a[0] + a[2]);
}

// ---- polynomial expansion -----------------------------------------------------------------------------------------
#define PE_TW 64
#define PE_TH 16
template <int PN>   // poly_n as a compile-time constant: the tile index divisions and the tap loops unroll
__global__ void __launch_bounds__(256) fb_polyexp_kernel(const float* __restrict__ I, float* __restrict__ R, int h, int w, FbPoly P) {
  __shared__ float src[PE_TH + 2 * PN][PE_TW + 2 * PN];
  __shared__ float row[3][PE_TH][PE_TW + 2 * PN];
  constexpr int n = PN;
  const int x0 = blockIdx.x * PE_TW, y0 = blockIdx.y * PE_TH;
  const float* img = I + (size_t)blockIdx.z * h * w;
  constexpr int sw = PE_TW + 2 * n, sh = PE_TH + 2 * n;
  for (int e = threadIdx.x; e < sw * sh; e += blockDim.x) {
    const int ly = e / sw, lx = e - ly * sw;
    const int gy = min(max(y0 + ly - n, 0), h - 1), gx = min(max(x0 + lx - n, 0), w - 1);  // replicate, as OpenCV's row buffer
    src[ly][lx] = img[(size_t)gy * w + gx];
  }
  __syncthreads();
  for (int e = threadIdx.x; e < PE_TH * sw; e += blockDim.x) {
    const int ly = e / sw, lx = e - ly * sw;
    float t0 = src[ly + n][lx] * P.g[0], t1 = 0.f, t2 = 0.f;
#pragma unroll
    for (int k = 1; k <= n; k++) {
      const float s0 = src[ly + n - k][lx], s1 = src[ly + n + k][lx], p = s0 + s1;
      t0 = t0 + P.g[k] * p;
      t1 = t1 + P.xg[k] * (s1 - s0);
      t2 = t2 + P.xxg[k] * p;
    }
    row[0][ly][lx] = t0; row[1][ly][lx] = t1; row[2][ly][lx] = t2;
  }
  __syncthreads();
  float* out = R + (size_t)blockIdx.z * 5 * h * w;
  const size_t plane = (size_t)h * w;
  for (int e = threadIdx.x; e < PE_TH * PE_TW; e += blockDim.x) {
    const int ly = e / PE_TW, lx = e - ly * PE_TW, x = lx + n;
    if (y0 + ly >= h || x0 + lx >= w) continue;
    const float g0 = P.g[0];
    double b1 = row[0][ly][x] * g0, b2 = 0, b3 = row[1][ly][x] * g0, b4 = 0, b5 = row[2][ly][x] * g0, b6 = 0;
#pragma unroll
    for (int k = 1; k <= n; k++) {
      const float p0 = row[0][ly][x + k], m0 = row[0][ly][x - k], p1 = row[1][ly][x + k], m1 = row[1][ly][x - k];
      const float p2 = row[2][ly][x + k], m2 = row[2][ly][x - k];
      const double tg = p0 + m0;
      const float gk = P.g[k];
      b1 += tg * gk;
      b4 += tg * P.xxg[k];
      b2 += (p0 - m0) * P.xg[k];
      b3 += (p1 + m1) * gk;
      b6 += (p1 - m1) * P.xg[k];
      b5 += (p2 + m2) * gk;
    }
    const size_t o = (size_t)(y0 + ly) * w + x0 + lx;
    out[o + plane] = (float)(b2 * P.ig11);
    out[o] = (float)(b3 * P.ig11);
    out[o + 3 * plane] = (float)(b1 * P.ig03 + b4 * P.ig33);
    out[o + 2 * plane] = (float)(b1 * P.ig03 + b5 * P.ig33);
    out[o + 4 * plane] = (float)(b6 * P.ig55);
  }
}

// ---- FarnebackUpdateMatrices for one pixel ------------------------------------------------------------------------
__device__ __forceinline__ void fb_matrix_core(const float* __restrict__ R1, size_t plane, int w, int h, int x, int y, float dx, float dy,
                                               float q0, float q1, float q2, float q3, float q4, float* __restrict__ m5) {
  float fx = (float)x + dx, fy = (float)y + dy;
  const int x1 = (int)floorf(fx), y1 = (int)floorf(fy);
  fx -= (float)x1; fy -= (float)y1;
  float r2, r3, r4, r5, r6;
  if ((unsigned)x1 < (unsigned)(w - 1) && (unsigned)y1 < (unsigned)(h - 1)) {
    const float a00 = (1.f - fx) * (1.f - fy), a01 = fx * (1.f - fy), a10 = (1.f - fx) * fy, a11 = fx * fy;
    const float* p = R1 + (size_t)y1 * w + x1;
    r2 = a00 * p[0] + a01 * p[1] + a10 * p[w] + a11 * p[w + 1];
    p += plane; r3 = a00 * p[0] + a01 * p[1] + a10 * p[w] + a11 * p[w + 1];
    p += plane; r4 = a00 * p[0] + a01 * p[1] + a10 * p[w] + a11 * p[w + 1];
    p += plane; r5 = a00 * p[0] + a01 * p[1] + a10 * p[w] + a11 * p[w + 1];
    p += plane; r6 = a00 * p[0] + a01 * p[1] + a10 * p[w] + a11 * p[w + 1];
    r4 = (q2 + r4) * 0.5f;
    r5 = (q3 + r5) * 0.5f;
    r6 = (q4 + r6) * 0.25f;
  } else {
    r2 = r3 = 0.f;
    r4 = q2; r5 = q3; r6 = q4 * 0.5f;
  }
  r2 = (q0 - r2) * 0.5f;
  r3 = (q1 - r3) * 0.5f;
  r2 += r4 * dy + r6 * dx;
  r3 += r6 * dy + r5 * dx;
  const int B = 5;
  if ((unsigned)(x - B) >= (unsigned)(w - B * 2) || (unsigned)(y - B) >= (unsigned)(h - B * 2)) {
    const float border[5] = {0.14f, 0.14f, 0.4472f, 0.4472f, 0.4472f};
    const float scale = (x < B ? border[x] : 1.f) * (x >= w - B ? border[w - x - 1] : 1.f) * (y < B ? border[y] : 1.f) *
                        (y >= h - B ? border[h - y - 1] : 1.f);
    r2 *= scale; r3 *= scale; r4 *= scale; r5 *= scale; r6 *= scale;
  }
  m5[0] = r4 * r4 + r6 * r6;
  m5[1] = (r4 + r5) * r6;
  m5[2] = r5 * r5 + r6 * r6;
  m5[3] = r4 * r2 + r6 * r3;
  m5[4] = r6 * r2 + r5 * r3;
}
__device__ __forceinline__ void fb_matrix_px(const float* __restrict__ R0, const float* __restrict__ R1, size_t plane, int w,
                                             int h, int x, int y, float dx, float dy, float* __restrict__ M) {
  const size_t o = (size_t)y * w + x;
  float m5[5];
  fb_matrix_core(R1, plane, w, h, x, y, dx, dy, R0[o], R0[o + plane], R0[o + 2 * plane], R0[o + 3 * plane], R0[o + 4 * plane], m5);
#pragma unroll
  for (int c = 0; c < 5; c++) M[o + c * plane] = m5[c];
}
// four adjacent pixels (x % 4 == 0, w % 4 == 0): the five coefficient planes of frame 0 come in and the five matrix planes go
// out as 16-byte accesses (10 instead of 40 LSU instructions; the 20 bilinear taps per pixel stay scalar gathers)
__device__ __forceinline__ void fb_matrix_px4(const float* __restrict__ R0, const float* __restrict__ R1, size_t plane, int w, int h,
                                              int x, int y, const float2* fl4, float* __restrict__ M) {
  const size_t o = (size_t)y * w + x;
  float4 q[5], m[5];
#pragma unroll
  for (int c = 0; c < 5; c++) q[c] = *reinterpret_cast<const float4*>(R0 + o + c * plane);
  float* mf = reinterpret_cast<float*>(m);
  const float* qf = reinterpret_cast<const float*>(q);
#pragma unroll
  for (int i = 0; i < 4; i++) {
    float m5[5];
    fb_matrix_core(R1, plane, w, h, x + i, y, fl4[i].x, fl4[i].y, qf[i], qf[4 + i], qf[8 + i], qf[12 + i], qf[16 + i], m5);
#pragma unroll
    for (int c = 0; c < 5; c++) mf[4 * c + i] = m5[c];
  }
#pragma unroll
  for (int c = 0; c < 5; c++) *reinterpret_cast<float4*>(M + o + c * plane) = m[c];
}


// first UpdateMatrices of a level: flow = resize(previous level's flow, INTER_LINEAR) * (1 / pyr_scale), zero at the coarsest level
__global__ void fb_matrices_kernel(const float* __restrict__ R, const float2* __restrict__ prev, int hp, int wp, float up,
                                   float* __restrict__ M, int h, int w, double sx, double sy) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, pair = blockIdx.z;
  if (x >= w) return;
  const size_t plane = (size_t)h * w;
  float dx = 0.f, dy = 0.f;
  if (prev) {
    const float2* pf = prev + (size_t)pair * hp * wp;
    const FTap tx = ftap(x, sx, wp, true), ty = ftap(y, sy, hp, false);   // sx, sy = 1 / (w / wp), 1 / (h / hp): computed once on the host
    const float2 a = pf[(size_t)ty.s0 * wp + tx.s0], b = pf[(size_t)ty.s0 * wp + tx.s1];
    const float2 c = pf[(size_t)ty.s1 * wp + tx.s0], d = pf[(size_t)ty.s1 * wp + tx.s1];
    const float h0x = a.x * tx.a0 + b.x * tx.a1, h1x = c.x * tx.a0 + d.x * tx.a1;
    const float h0y = a.y * tx.a0 + b.y * tx.a1, h1y = c.y * tx.a0 + d.y * tx.a1;
    dx = (h0x * ty.a0 + h1x * ty.a1) * up;
    dy = (h0y * ty.a0 + h1y * ty.a1) * up;
  }
  const float* R0 = R + (size_t)(2 * pair) * 5 * plane;
  fb_matrix_px(R0, R0 + 5 * plane, plane, w, h, x, y, dx, dy, M + (size_t)pair * 5 * plane);
}

// ---- box blur of M + 2x2 solve (+ next UpdateMatrices) --------------------------------------------------------------
// A CTA owns a strip of BS_COLS output columns and a segment of rows and walks DOWN it (OpenCV's own structure,
// FarnebackUpdateFlow_Blur: vertical running sums per column, then horizontal running sums, both in double):
//   V phase  thread = column of the strip + its 2m halo columns (clamped = OpenCV's replicated border): the five fp64
//            vertical window sums are kept in registers and updated per row with the entering / leaving row of M (read
//            from global memory: the leaving row is an L1/L2 hit, it entered 2m+1 rows earlier); BS_R rows are written to
//            a shared-memory tile per block barrier;
//   H phase  thread = (row, run of 8 columns): horizontal sliding sums over the tile (positions skewed by one double per
//            8 so that the 16 runs of a row hit distinct banks), the 2x2 solve, flow to shared memory;
//   S phase  thread = column again: coalesced flow store (last iteration) or the next UpdateMatrices into Mout.
// The row segmentation depends only on the level's height, never on the batch: a pair's flow is bit-identical alone or
// in a batch.
#define BS_COLS 128
#define BS_THREADS (BS_COLS + 2 * FB_MAX_M + 2)            // 160
#define BS_POS(x) ((x) + ((x) >> (RUN == 8 ? 3 : 2)))      // one double of skew per run
#define BS_STRIDE (BS_THREADS + BS_THREADS / RUN)          // doubles per (row, plane)
template <int BS_R, int RUN, bool FUSE>
__global__ void __launch_bounds__(BS_THREADS) fb_blur_solve_kernel(const float* __restrict__ Min, const float* __restrict__ R,
                                                                   float* __restrict__ Mout /*null on the last iteration*/,
                                                                   float2* __restrict__ flow /*null unless last iteration*/, int h, int w, int m,
                                                                   int rows_per_seg) {
  extern __shared__ double tile[];  // [BS_R][5][BS_STRIDE]
  __shared__ float2 fl[BS_R][BS_COLS];
  const int t = threadIdx.x, pair = blockIdx.z, x0 = blockIdx.x * BS_COLS;
  const int y0 = blockIdx.y * rows_per_seg, y1 = min(h, y0 + rows_per_seg);
  const int win = 2 * m + 1;
  const size_t plane = (size_t)h * w;
  const float* Mi = Min + (size_t)pair * 5 * plane;
  const bool vcol = t < BS_COLS + 2 * m;
  const int gx = min(max(x0 + t - m, 0), w - 1);
  double S[5] = {0.0, 0.0, 0.0, 0.0, 0.0};
  if (vcol) {
    for (int j = -m; j <= m; j++) {
      const size_t o = (size_t)min(max(y0 + j, 0), h - 1) * w + gx;
#pragma unroll
      for (int c = 0; c < 5; c++) S[c] += (double)Mi[c * plane + o];
    }
  }
  const int tp = BS_POS(t);
  constexpr int RUNS = BS_COLS / RUN;
  const int hr = t / RUNS, hx = (t % RUNS) * RUN;          // H-phase role (threads 0 .. BS_R * RUNS - 1)
  const float* R0 = R + (size_t)(2 * pair) * 5 * plane;
  const double scale = 1.0 / ((double)win * (double)win);
  for (int yb = y0; yb < y1; yb += BS_R) {
    const int nr = min(BS_R, y1 - yb);
    if (vcol) {
      float en[BS_R][5], le[BS_R][5];
#pragma unroll
      for (int r = 0; r < BS_R; r++) {                     // the batch's loads first: 2 x 5 x BS_R independent requests
        const size_t oe = (size_t)min(yb + r + m + 1, h - 1) * w + gx, ol = (size_t)max(yb + r - m, 0) * w + gx;
#pragma unroll
        for (int c = 0; c < 5; c++) {
          en[r][c] = r < nr ? Mi[c * plane + oe] : 0.f;
          le[r][c] = r < nr ? Mi[c * plane + ol] : 0.f;
        }
      }
#pragma unroll
      for (int r = 0; r < BS_R; r++) {
#pragma unroll
        for (int c = 0; c < 5; c++) {
          tile[(r * 5 + c) * BS_STRIDE + tp] = S[c];
          S[c] += (double)en[r][c] - (double)le[r][c];
        }
      }
    }
    __syncthreads();
    if (t < BS_R * RUNS && hr < nr) {
      double g[5];
      const double* base = tile + (size_t)hr * 5 * BS_STRIDE;
#pragma unroll
      for (int c = 0; c < 5; c++) {
        const double* rowp = base + c * BS_STRIDE;
        double a = 0.0;
        for (int j = 0; j < win; j++) a += rowp[BS_POS(hx + j)];
        g[c] = a;
      }
#pragma unroll
      for (int i = 0; i < RUN; i++) {
        const double g11 = g[0] * scale, g12 = g[1] * scale, g22 = g[2] * scale, h1 = g[3] * scale, h2 = g[4] * scale;
        const double idet = 1.0 / (g11 * g22 - g12 * g12 + 1e-3);
        fl[hr][hx + i] = make_float2((float)((g11 * h2 - g12 * h1) * idet), (float)((g22 * h1 - g12 * h2) * idet));
        if (i < RUN - 1) {
#pragma unroll
          for (int c = 0; c < 5; c++) {
            const double* rowp = base + c * BS_STRIDE;
            g[c] += rowp[BS_POS(hx + i + win)] - rowp[BS_POS(hx + i)];
          }
        }
      }
    }
    __syncthreads();
    if (t < BS_COLS && x0 + t < w) {
      const int x = x0 + t;
      for (int r = 0; r < nr; r++) {
        const int y = yb + r;
        const float2 f = fl[r][t];
        if (flow) flow[(size_t)pair * plane + (size_t)y * w + x] = f;
        if (FUSE && Mout) fb_matrix_px(R0, R0 + 5 * plane, plane, w, h, x, y, f.x, f.y, Mout + (size_t)pair * 5 * plane);
      }
    }
  }
}
static int bs_rows_per_seg(int h) { return h >= 1024 ? 128 : h >= 512 ? 64 : 32; }   // a function of the level only
// UpdateMatrices from a flow field of the level's own size (the unfused variant of the blur kernel's S phase)
__global__ void fb_matrices_flow_kernel(const float* __restrict__ R, const float2* __restrict__ flow, float* __restrict__ M, int h, int w) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, pair = blockIdx.z;
  if (x >= w) return;
  const size_t plane = (size_t)h * w;
  const float2 f = flow[(size_t)pair * plane + (size_t)y * w + x];
  const float* R0 = R + (size_t)(2 * pair) * 5 * plane;
  fb_matrix_px(R0, R0 + 5 * plane, plane, w, h, x, y, f.x, f.y, M + (size_t)pair * 5 * plane);
}
__global__ void fb_matrices_flow4_kernel(const float* __restrict__ R, const float2* __restrict__ flow, float* __restrict__ M, int h, int w) {
  const int x = (blockIdx.x * blockDim.x + threadIdx.x) * 4, y = blockIdx.y, pair = blockIdx.z;
  if (x >= w) return;
  const size_t plane = (size_t)h * w;
  const float4* fp = reinterpret_cast<const float4*>(flow + (size_t)pair * plane + (size_t)y * w + x);
  const float4 fa = fp[0], fb = fp[1];
  const float2 fl4[4] = {make_float2(fa.x, fa.y), make_float2(fa.z, fa.w), make_float2(fb.x, fb.y), make_float2(fb.z, fb.w)};
  const float* R0 = R + (size_t)(2 * pair) * 5 * plane;
  fb_matrix_px4(R0, R0 + 5 * plane, plane, w, h, x, y, fl4, M + (size_t)pair * 5 * plane);
}
static int fb_variant() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("VFI_FB_VARIANT"); v = e ? atoi(e) : 0; }
  return v;
}
template <int BS_R, int RUN, bool FUSE>
static void bs_launch(dim3 grid, cudaStream_t s, const float* Mi, const float* R, float* Mo, float2* fl, int hh, int w, int m, int rps) {
  const size_t smem = (size_t)BS_R * 5 * (BS_THREADS + BS_THREADS / RUN) * sizeof(double);
  static bool set = false;
  if (!set) { cudaFuncSetAttribute(fb_blur_solve_kernel<BS_R, RUN, FUSE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); set = true; }
  fb_blur_solve_kernel<BS_R, RUN, FUSE><<<grid, BS_THREADS, smem, s>>>(Mi, R, Mo, fl, hh, w, m, rps);
}

int cv_round(double v) { return (int)nearbyint(v); }

// cv::getGaussianKernel(n, sigma, CV_32F)
void gaussian_kernel(int n, double sigma, FbKern& kk) {
  kk.r = n / 2;
  if (sigma <= 0 && n == 3) { kk.k[0] = 0.5f; kk.k[1] = 0.25f; return; }
  const double s = sigma > 0 ? sigma : ((n - 1) * 0.5 - 1) * 0.3 + 0.8;
  std::vector<double> v(n);
  double sum = 0;
  for (int i = 0; i < n; i++) { const double x = i - (n - 1) * 0.5; v[i] = exp(-0.5 / (s * s) * x * x); sum += v[i]; }
  for (int d = 0; d <= kk.r; d++) kk.k[d] = (float)(v[kk.r + d] / sum);
}

// 6x6 symmetric positive-definite inverse by Gauss-Jordan (FarnebackPrepareGaussian's G.inv(DECOMP_CHOLESKY))
void invert6(double G[6][6], double inv[6][6]) {
  double a[6][12];
  for (int i = 0; i < 6; i++)
    for (int j = 0; j < 6; j++) { a[i][j] = G[i][j]; a[i][6 + j] = i == j; }
  for (int c = 0; c < 6; c++) {
    int p = c;
    for (int r = c + 1; r < 6; r++) if (fabs(a[r][c]) > fabs(a[p][c])) p = r;
    if (p != c) for (int j = 0; j < 12; j++) std::swap(a[p][j], a[c][j]);
    const double d = a[c][c];
    for (int j = 0; j < 12; j++) a[c][j] /= d;
    for (int r = 0; r < 6; r++) if (r != c) { const double f = a[r][c]; if (f != 0) for (int j = 0; j < 12; j++) a[r][j] -= f * a[c][j]; }
  }
  for (int i = 0; i < 6; i++) for (int j = 0; j < 6; j++) inv[i][j] = a[i][6 + j];
}

void prepare_poly(int n, double sigma, FbPoly& P) {
  if (sigma < 1.1920928955078125e-07) sigma = n * 0.3;
  float g[2 * FB_MAX_N + 1], *gc = g + n;
  double s = 0;
  for (int x = -n; x <= n; x++) { gc[x] = (float)exp(-x * x / (2 * sigma * sigma)); s += gc[x]; }
  s = 1. / s;
  P.n = n;
  for (int x = -n; x <= n; x++) gc[x] = (float)(gc[x] * s);
  for (int x = 0; x <= n; x++) { P.g[x] = gc[x]; P.xg[x] = (float)(x * gc[x]); P.xxg[x] = (float)(x * x * gc[x]); }
  double G[6][6] = {{0}}, inv[6][6];
  for (int y = -n; y <= n; y++)
    for (int x = -n; x <= n; x++) {
      const float wgt = gc[y] * gc[x];
      G[0][0] += wgt;
      G[1][1] += wgt * x * x;
      G[3][3] += wgt * x * x * x * x;
      G[5][5] += wgt * x * x * y * y;
    }
  G[2][2] = G[0][3] = G[0][4] = G[3][0] = G[4][0] = G[1][1];
  G[4][4] = G[3][3];
  G[3][4] = G[4][3] = G[5][5];
  invert6(G, inv);
  P.ig11 = inv[1][1]; P.ig03 = inv[0][3]; P.ig33 = inv[3][3]; P.ig55 = inv[5][5];
}

}  // namespace

extern "C" __attribute__((visibility("default"))) int vfi_farneback(vfi_handle* h, const uint8_t* img0, const uint8_t* img1, int n, int H, int W,
                                                                      int channels, double pyr_scale, int levels, int winsize, int iterations,
                                                                      int poly_n, double poly_sigma, float* flow_out, double* stats_out,
                                                                      void* stream) {
  if (!h) return VFI_E_ARG;
  if (!img0 || !img1 || n < 1 || H < 1 || W < 1 || (channels != 1 && channels != 3) || (!flow_out && !stats_out))
    VFI_FAIL(h, VFI_E_ARG, "bad argument");
  if (!(pyr_scale > 0.0 && pyr_scale < 1.0) || levels < 0 || levels > 16 || iterations < 1 || (poly_n != 5 && poly_n != 7) ||
      winsize < 1 || winsize / 2 > FB_MAX_M)
    VFI_FAIL(h, VFI_E_ARG, "unsupported Farneback parameters (0 < pyr_scale < 1, poly_n 5 or 7, winsize <= %d)", 2 * FB_MAX_M + 1);
  VFI_ENTER(h, stream);
  cudaStream_t s = (cudaStream_t)stream;
  // number of pyramid levels actually used (optflowgf.cpp: stop when a side would drop under 32 px)
  int nlev = 0;
  {
    double sc = 1.0;
    for (; nlev < levels; nlev++) { sc *= pyr_scale; if (W * sc < 32 || H * sc < 32) break; }
  }
  FbPoly P;
  prepare_poly(poly_n, poly_sigma, P);
  const int m = winsize / 2;
  const size_t px = (size_t)H * W;
  const int chunk = std::min(n, 8);
  Carve cv{nullptr, 0};
  std::vector<unsigned long long> guard_gaps;              // vfi_set_option("ws_guard"): pattern gaps around every buffer, checked below
  if (h->opt_ws_guard) { cv.guard = WS_GUARD_BYTES; cv.gaps = &guard_gaps; }
  for (int pass = 0; pass < 2; pass++) {
    cv.off = 0;
    cv.base = pass ? (uint8_t*)h->ws : nullptr;
    if (cv.guard) cv.take<uint8_t>(0);
    uint8_t* gray = cv.take<uint8_t>((size_t)chunk * 2 * px);
    float* A = cv.take<float>((size_t)chunk * 2 * px);
    float* I = cv.take<float>((size_t)chunk * 2 * px);
    float* R = cv.take<float>((size_t)chunk * 2 * 5 * px);
    float* M0 = cv.take<float>((size_t)chunk * 5 * px);
    float* M1 = cv.take<float>((size_t)chunk * 5 * px);
    float2* fA = cv.take<float2>((size_t)chunk * px);
    float2* fB = cv.take<float2>((size_t)chunk * px);
    float2* tmpf = cv.take<float2>((size_t)chunk * px);
    const size_t guarded_bytes = cv.off;
    unsigned long long* guard_dev = cv.guard ? ws_guard_scratch(cv) : nullptr;
    if (pass == 0) { if (int rc = vfi_ws_reserve(h, cv.off + 4096)) return rc; continue; }
    if (cv.guard) VFI_CUDA(h, cudaMemsetAsync(h->ws, WS_GUARD_BYTE, guarded_bytes, s));
    for (int n0 = 0; n0 < n; n0 += chunk) {
      const int nn = std::min(chunk, n - n0);
      ProfScope ps(h, PC_REDUCE, s);
      fb_gray_kernel<<<dim3(cdiv(px, 256 * 8), 2 * nn), 256, 0, s>>>(img0 + (size_t)n0 * px * channels, img1 + (size_t)n0 * px * channels, px,
                                                                   channels, gray);
      VFI_LAUNCHED(h);
      float2* prev = nullptr;
      int hp = 0, wp = 0;
      for (int k = nlev; k >= 0; k--) {
        double scale = 1.0;
        for (int i = 0; i < k; i++) scale *= pyr_scale;
        const double sigma = (1. / scale - 1) * 0.5;
        int smooth = cv_round(sigma * 5) | 1;
        smooth = std::max(smooth, 3);
        if (smooth / 2 > FB_MAX_KR) VFI_FAIL(h, VFI_E_ARG, "pyramid too deep for this build (kernel radius %d)", smooth / 2);
        const int w = cv_round(W * scale), hh = cv_round(H * scale);
        FbKern kk;
        gaussian_kernel(smooth, sigma, kk);
        if (w == W && hh == H && kk.r == 1 && fb_variant() != 9) {
          fb_blur3_kernel<<<dim3(cdiv(W, 128), H, 2 * nn), 128, 0, s>>>(gray, I, H, W, kk.k[0], kk.k[1]);
          VFI_LAUNCHED(h);
        } else {
          fb_blur_h_kernel<<<dim3(cdiv(w, 128), H, 2 * nn), 128, 0, s>>>(gray, A, H, W, w, 1.0 / ((double)w / (double)W), kk);
          VFI_LAUNCHED(h);
          fb_blur_v_kernel<<<dim3(cdiv(w, 128), hh, 2 * nn), 128, 0, s>>>(A, I, H, hh, w, 1.0 / ((double)hh / (double)H), kk);
          VFI_LAUNCHED(h);
        }
        if (poly_n == 5) fb_polyexp_kernel<5><<<dim3(cdiv(w, PE_TW), cdiv(hh, PE_TH), 2 * nn), 256, 0, s>>>(I, R, hh, w, P);
        else fb_polyexp_kernel<7><<<dim3(cdiv(w, PE_TW), cdiv(hh, PE_TH), 2 * nn), 256, 0, s>>>(I, R, hh, w, P);
        VFI_LAUNCHED(h);
        fb_matrices_kernel<<<dim3(cdiv(w, 128), hh, nn), 128, 0, s>>>(R, prev, hp, wp, (float)(1. / pyr_scale), M0, hh, w,
                                                                      prev ? 1.0 / ((double)w / (double)wp) : 1.0, prev ? 1.0 / ((double)hh / (double)hp) : 1.0);
        VFI_LAUNCHED(h);
        float* Mi = M0;
        float* Mo = M1;
        float2* cur = (k == 0 && flow_out) ? reinterpret_cast<float2*>(flow_out) + (size_t)n0 * px : (prev == fA ? fB : fA);
        for (int i = 0; i < iterations; i++) {
          const bool last = i == iterations - 1;
          const int rps = bs_rows_per_seg(hh);
          const dim3 grid(cdiv(w, BS_COLS), cdiv(hh, rps), nn);
          const int var = fb_variant();
          if (var == 0 || var == 1) {
            // shipped: blur + solve (flow only), then UpdateMatrices as its own full-occupancy launch -- fused behind the
            // solve it ran at the blur kernel's low occupancy and cost more than both launches together (A/B on B200:
            // 0.71 ms per 1080p pair fused, 0.57 unfused)
            if (var == 0) bs_launch<4, 8, false>(grid, s, Mi, R, nullptr, last ? cur : tmpf, hh, w, m, rps);
            else bs_launch<8, 8, false>(grid, s, Mi, R, nullptr, last ? cur : tmpf, hh, w, m, rps);
            if (!last) {
              VFI_LAUNCHED(h);
              if ((w & 3) == 0 && var == 0) fb_matrices_flow4_kernel<<<dim3(cdiv(w / 4, 64), hh, nn), 64, 0, s>>>(R, tmpf, Mo, hh, w);
              else fb_matrices_flow_kernel<<<dim3(cdiv(w, 128), hh, nn), 128, 0, s>>>(R, tmpf, Mo, hh, w);
            }
          } else if (var == 2) bs_launch<4, 4, true>(grid, s, Mi, R, last ? nullptr : Mo, last ? cur : nullptr, hh, w, m, rps);
          else if (var == 3) bs_launch<4, 8, true>(grid, s, Mi, R, last ? nullptr : Mo, last ? cur : nullptr, hh, w, m, rps);
          else bs_launch<8, 8, true>(grid, s, Mi, R, last ? nullptr : Mo, last ? cur : nullptr, hh, w, m, rps);
          VFI_LAUNCHED(h);
          std::swap(Mi, Mo);
        }
        prev = cur; hp = hh; wp = w;
      }
      if (stats_out)
        if (int rc = vfi_flow_stats_impl(h, reinterpret_cast<const float*>(prev), nn, H, W, stats_out + (size_t)3 * n0, s)) return rc;
    }
    if (cv.guard)
      if (int rc = vfi_ws_guard_check(h, guard_gaps, guard_dev, s)) return rc;
  }
  return VFI_OK;
}
