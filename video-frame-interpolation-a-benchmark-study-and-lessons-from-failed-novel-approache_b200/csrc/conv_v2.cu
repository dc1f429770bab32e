// conv_v2.cu -- second-generation implicit-GEMM 3x3 convolution for the layers that dominate the step
// (IFBlock ResConv chains of blocks 3/4, Head cnn1/cnn2): weights RESIDENT in shared memory, the input
// halo tile loaded ONCE per tile (not once per tap), all nine taps issued as tcgen05.mma from the same
// buffer through shifted no-swizzle K-major descriptors.
//
//   macro tile   : 16 rows x (8*KW) columns of output pixels = KW MMA tiles of M = 128 (16 rows x 8 px)
//   A (smem)     : [Cin/8 chunks][18 x (8*KW+2) halo pixels][8 ch] bf16.  In the canonical no-swizzle K-major
//                  layout ((8,n),2):((1,SBO),LBO) (16-byte units) row r of the MMA sits at
//                  start + (r%8)*16 + (r/8)*SBO; with SBO = halo row pitch, the 16 groups of 8 rows are the 16
//                  image rows of an 8-px-wide tile, and tap (dy,dx) is just start += (dy*pitch + dx)*16.
//   B (smem)     : [tap][Cin/8][Cout][8] bf16, copied once per CTA (persistent kernel), SBO = 128 B, LBO = Cout*16.
//   D (TMEM)     : KW accumulators of Cout fp32 columns, double buffered (2*KW*Cout <= 512 columns).
//   epilogue     : tcgen05.ld -> +bias -> (*beta + residual, taken from the SAME smem halo tile) -> LeakyReLU
//                  -> bf16 NHWC store.
// Warp roles (288 threads): warps 0-3 producers (LDG.128 -> STS.128, zero fill = conv padding), warp 4 TMEM
// allocator + single-thread MMA issuer, warps 5-8 epilogue.
// Layer semantics: SURVEY.md Appendix A.2 (conv(), ResConv).
#include <algorithm>

#include "common.cuh"

struct V2Params {
  const bf16* in;
  bf16* out;
  const bf16* wpack;
  const float* bias;
  const float* beta;  // null => plain conv
  int N, H, W, Cin, Cout;
  int KW, tiles_x, tiles_y, num_tiles;
  int pitch;   // halo row pitch in pixels = 8*KW + 2
  int npix;    // 18 * pitch
  int CS;      // chunk stride in 16-byte units (>= npix, padded against bank conflicts)
  int a_bytes, nstages, w_bytes, tmem_cols, acc_stride;
  int act, chunk_shift;  // chunk_shift = log2(Cin/8) or -1
  float slope;
};

#define V2_THREADS 288
#define V2_PRODUCERS 128
#define V2_MAX_STAGES 4
#define V2_UNROLL 11

namespace {
__device__ __forceinline__ uint32_t s_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void bar_init(uint64_t* b, uint32_t cnt) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s_u32(b)), "r"(cnt) : "memory");
}
__device__ __forceinline__ void bar_wait(uint64_t* b, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred P1;\n\t"
      "V2_WAIT:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t"
      "@P1 bra V2_DONE;\n\t"
      "bra V2_WAIT;\n\t"
      "V2_DONE:\n\t}" ::"r"(s_u32(b)), "r"(parity)
      : "memory");
}
__device__ __forceinline__ void bar_arrive(uint64_t* b) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(s_u32(b)) : "memory");
}
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void commit_to(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(s_u32(bar))
               : "memory");
}
__device__ __forceinline__ void mma_bf16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                         uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void ld16(uint32_t taddr, float* v) {
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 16; i++) v[i] = __uint_as_float(r[i]);
}
// no-swizzle K-major descriptor: start, LBO (K-direction core-matrix stride), SBO (8-row group stride), bytes
__device__ __forceinline__ uint64_t desc_ns(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = (uint64_t)((saddr & 0x3FFFF) >> 4);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;  // descriptor version (sm_100)
  return d;                // layout_type 0 = SWIZZLE_NONE, base_offset 0
}
__device__ __forceinline__ uint32_t idesc_bf16(int n) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | (8u << 24);  // F32 acc, BF16 x BF16, M = 128
}
__device__ __forceinline__ uint32_t pk_bf16(float a, float b) {
  __nv_bfloat162 t = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&t);
}
}  // namespace

__global__ void __launch_bounds__(V2_THREADS, 1) conv_v2_kernel(const __grid_constant__ V2Params P) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint64_t full_bar[V2_MAX_STAGES], empty_bar[V2_MAX_STAGES], tfull_bar[2], tempty_bar[2];
  __shared__ uint32_t tmem_base_smem;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint8_t* w_s = smem;
  float* bias_s = reinterpret_cast<float*>(smem + P.w_bytes);
  float* beta_s = bias_s + P.Cout;
  uint8_t* a_s = smem + P.w_bytes + ((2 * P.Cout * 4 + 127) & ~127);
  const bool use_res = P.beta != nullptr;

  // ---- one-time setup: barriers, TMEM, resident weights ----
  if (threadIdx.x == 0) {
    for (int i = 0; i < P.nstages; i++) { bar_init(&full_bar[i], V2_PRODUCERS); bar_init(&empty_bar[i], use_res ? 5 : 1); }
    for (int i = 0; i < 2; i++) { bar_init(&tfull_bar[i], 1); bar_init(&tempty_bar[i], 4); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 4) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s_u32(&tmem_base_smem)),
                 "r"((uint32_t)P.tmem_cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  {
    const uint4* src = reinterpret_cast<const uint4*>(P.wpack);
    uint4* dst = reinterpret_cast<uint4*>(w_s);
    for (int i = threadIdx.x; i < P.w_bytes / 16; i += V2_THREADS) dst[i] = __ldg(src + i);
    for (int i = threadIdx.x; i < P.Cout; i += V2_THREADS) {
      bias_s[i] = P.bias[i];
      beta_s[i] = use_res ? P.beta[i] : 1.f;
    }
    fence_async_smem();
  }
  tc_before();
  __syncthreads();
  tc_after();
  const uint32_t tmem_base = tmem_base_smem;
  const int nchunk = P.Cin >> 3;

  if (warp < 4) {
    // ================= producers: halo tile -> smem [chunk][pixel][8ch] =================
    int it = 0;
    for (int tile = blockIdx.x; tile < P.num_tiles; tile += gridDim.x, it++) {
      const int st = it % P.nstages;
      const uint32_t ph = (uint32_t)((it / P.nstages) & 1);
      const int tx = tile % P.tiles_x, ty = (tile / P.tiles_x) % P.tiles_y, n = tile / (P.tiles_x * P.tiles_y);
      const int gx0 = tx * 8 * P.KW - 1, gy0 = ty * 16 - 1;
      bar_wait(&empty_bar[st], ph ^ 1);
      uint4* dst = reinterpret_cast<uint4*>(a_s + (size_t)st * P.a_bytes);
      const bf16* img = P.in + (size_t)n * P.H * P.W * P.Cin;
      // flattened (halo row, pixel, chunk) items, V2_UNROLL independent 16-byte loads in flight per thread
      const int row_items = P.pitch * nchunk, total = 18 * row_items;
      int py = 0, rem = threadIdx.x;  // item = py * row_items + rem
      while (rem >= row_items) { rem -= row_items; py++; }
      for (int i0 = threadIdx.x; i0 < total; i0 += V2_PRODUCERS * V2_UNROLL) {
        uint4 v[V2_UNROLL];
        int off[V2_UNROLL];
#pragma unroll
        for (int u = 0; u < V2_UNROLL; u++) {
          const int px = P.chunk_shift >= 0 ? (rem >> P.chunk_shift) : rem / nchunk;
          const int j = rem - px * nchunk;
          const int gy = gy0 + py, gx = gx0 + px;
          const bool ok = (i0 + u * V2_PRODUCERS < total) && gy >= 0 && gy < P.H && gx >= 0 && gx < P.W;
          off[u] = (i0 + u * V2_PRODUCERS < total) ? j * P.CS + py * P.pitch + px : -1;
          v[u] = make_uint4(0, 0, 0, 0);
          if (ok) v[u] = __ldg(reinterpret_cast<const uint4*>(img + ((size_t)gy * P.W + gx) * P.Cin + j * 8));
          rem += V2_PRODUCERS;
          while (rem >= row_items) { rem -= row_items; py++; }
        }
#pragma unroll
        for (int u = 0; u < V2_UNROLL; u++)
          if (off[u] >= 0) dst[off[u]] = v[u];
      }
      fence_async_smem();
      bar_arrive(&full_bar[st]);
    }
  } else if (warp == 4) {
    // ================= MMA issuer =================
    if (lane == 0) {
      const uint32_t idesc = idesc_bf16(P.Cout);
      const uint32_t a_lbo = (uint32_t)P.CS * 16, a_sbo = (uint32_t)P.pitch * 16;
      const uint32_t b_lbo = (uint32_t)P.Cout * 16, b_sbo = 128;
      const uint32_t w_addr = s_u32(w_s);
      const int ksteps = P.Cin >> 4;
      int it = 0;
      for (int tile = blockIdx.x; tile < P.num_tiles; tile += gridDim.x, it++) {
        const int st = it % P.nstages;
        const uint32_t ph = (uint32_t)((it / P.nstages) & 1);
        const int as = it & 1;
        bar_wait(&tempty_bar[as], (uint32_t)(((it >> 1) & 1) ^ 1));
        bar_wait(&full_bar[st], ph);
        tc_after();
        const uint32_t a_addr = s_u32(a_s + (size_t)st * P.a_bytes);
        // consecutive MMAs go to DIFFERENT accumulators (j innermost): back-to-back MMAs into the same TMEM columns
        // serialise on the accumulate dependency (~200 cycles each at these small N), independent ones pipeline.
        const uint32_t dcol0 = tmem_base + (uint32_t)(as * P.acc_stride);
        for (int tap = 0; tap < 9; tap++) {
          const int r = tap / 3, q = tap - 3 * r;
          const uint32_t a0 = a_addr + (uint32_t)(r * P.pitch + q) * 16;
          const uint32_t b0 = w_addr + (uint32_t)(tap * nchunk) * b_lbo;
          for (int k = 0; k < ksteps; k++) {
            const uint64_t bd = desc_ns(b0 + 2 * k * b_lbo, b_lbo, b_sbo);
            const uint64_t ad = desc_ns(a0 + 2 * k * a_lbo, a_lbo, a_sbo);
            const uint32_t acc = (tap | k) ? 1u : 0u;
            for (int j = 0; j < P.KW; j++)
              mma_bf16(dcol0 + (uint32_t)(j * P.Cout), ad + (uint64_t)(8 * j), bd, idesc, acc);  // +8 px = +128 B = +8 (16-B units)
          }
        }
        commit_to(&empty_bar[st]);
        commit_to(&tfull_bar[as]);
      }
    }
  } else {
    // ================= epilogue =================
    const int lg = warp & 3;
    const int m = lg * 32 + lane;
    const int ly = m >> 3, lx = m & 7;
    int it = 0;
    for (int tile = blockIdx.x; tile < P.num_tiles; tile += gridDim.x, it++) {
      const int st = it % P.nstages;
      const int as = it & 1;
      const int tx = tile % P.tiles_x, ty = (tile / P.tiles_x) % P.tiles_y, n = tile / (P.tiles_x * P.tiles_y);
      const int y = ty * 16 + ly;
      bar_wait(&tfull_bar[as], (uint32_t)((it >> 1) & 1));
      tc_after();
      const uint4* a_st = reinterpret_cast<const uint4*>(a_s + (size_t)st * P.a_bytes);
      for (int j = 0; j < P.KW; j++) {
        const int x = (tx * P.KW + j) * 8 + lx;
        const bool valid = (y < P.H) && (x < P.W);
        const uint32_t taddr = tmem_base + ((uint32_t)(lg * 32) << 16) + (uint32_t)(as * P.acc_stride + j * P.Cout);
        bf16* op = P.out + (((size_t)n * P.H + y) * P.W + x) * P.Cout;
        const int pc = (ly + 1) * P.pitch + 8 * j + lx + 1;  // centre pixel of this thread in the halo tile
        for (int c = 0; c < P.Cout; c += 16) {
          float v[16];
          ld16(taddr + c, v);
          if (valid) {
#pragma unroll
            for (int i = 0; i < 16; i++) v[i] += bias_s[c + i];
            if (use_res) {
              const uint4 r0 = a_st[(c >> 3) * P.CS + pc], r1 = a_st[((c >> 3) + 1) * P.CS + pc];
              const uint32_t rr[8] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
#pragma unroll
              for (int i = 0; i < 8; i++) {
                float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&rr[i]));
                v[2 * i] = v[2 * i] * beta_s[c + 2 * i] + f.x;
                v[2 * i + 1] = v[2 * i + 1] * beta_s[c + 2 * i + 1] + f.y;
              }
            }
            if (P.act) {
#pragma unroll
              for (int i = 0; i < 16; i++) v[i] = v[i] > 0.f ? v[i] : v[i] * P.slope;
            }
            uint4 o0, o1;
            o0.x = pk_bf16(v[0], v[1]); o0.y = pk_bf16(v[2], v[3]); o0.z = pk_bf16(v[4], v[5]); o0.w = pk_bf16(v[6], v[7]);
            o1.x = pk_bf16(v[8], v[9]); o1.y = pk_bf16(v[10], v[11]); o1.z = pk_bf16(v[12], v[13]); o1.w = pk_bf16(v[14], v[15]);
            *reinterpret_cast<uint4*>(op + c) = o0;
            *reinterpret_cast<uint4*>(op + c + 8) = o1;
          }
        }
      }
      tc_before();
      __syncwarp();
      if (lane == 0) {
        bar_arrive(&tempty_bar[as]);
        if (use_res) bar_arrive(&empty_bar[st]);
      }
    }
  }
  tc_before();
  __syncthreads();
  if (warp == 4) {
    tc_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)P.tmem_cols)
                 : "memory");
  }
}

// =============================================================================================
static uint16_t f2bf_v2(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffff) > 0x7f800000) return 0x7fc0;
  u += 0x7fff + ((u >> 16) & 1);
  return (uint16_t)(u >> 16);
}

bool v2_supported(const Layer& L) {
  if (L.kind != LK_CONV3 || L.stride != 1) return false;
  if (L.cin % 16 || L.cout % 16 || L.cin != L.cin_pad) return false;
  if (L.has_beta && L.cin != L.cout) return false;
  return (size_t)9 * L.cin * L.cout * 2 <= 80 * 1024;  // weights must stay resident next to >= 2 A stages
}

// Packs Conv2d weights [Co][Ci][3][3] into the resident smem image [tap][Ci/8][Co][8] bf16.
int v2_build_layer(vfi_handle* h, Layer& L, const std::vector<float>& w) {
  if (!v2_supported(L)) return 0;
  const int nch = L.cin / 8;
  std::vector<uint16_t> pk((size_t)9 * L.cin * L.cout);
  for (int tap = 0; tap < 9; tap++)
    for (int j = 0; j < nch; j++)
      for (int co = 0; co < L.cout; co++)
        for (int e = 0; e < 8; e++) {
          const int ci = j * 8 + e;
          pk[(((size_t)tap * nch + j) * L.cout + co) * 8 + e] = f2bf_v2(w[((size_t)co * L.cin + ci) * 9 + tap]);
        }
  VFI_CUDA(h, cudaMalloc(&L.wv2, pk.size() * 2));
  VFI_CUDA(h, cudaMemcpy(L.wv2, pk.data(), pk.size() * 2, cudaMemcpyHostToDevice));
  return 0;
}

int k_conv_v2(vfi_handle* h, const Layer& L, const bf16* in, bf16* out, int N, int H, int W, float slope,
              cudaStream_t s) {
  ProfScope _ps(h, PC_CONV_TC, s);
  if (!L.wv2) VFI_FAIL(h, VFI_E_STATE, "conv_v2: layer not packed");
  V2Params P;
  memset(&P, 0, sizeof(P));
  P.in = in; P.out = out; P.wpack = L.wv2; P.bias = L.bias; P.beta = L.has_beta ? L.beta : nullptr;
  P.N = N; P.H = H; P.W = W; P.Cin = L.cin; P.Cout = L.cout;
  P.act = L.act ? 1 : 0;
  P.slope = slope;
  P.w_bytes = 9 * L.cin * L.cout * 2;
  const int nchunk = L.cin / 8;
  P.chunk_shift = -1;
  for (int b = 0; b < 6; b++) if ((1 << b) == nchunk) P.chunk_shift = b;
  const size_t budget = 220 * 1024 - P.w_bytes - 1024;
  int kw = 4;
  for (; kw >= 1; kw >>= 1) {
    if (2 * kw * L.cout > 512) continue;
    if (kw > 1 && W <= 8 * (kw / 2)) continue;  // do not tile wider than the image needs
    const int pitch = 8 * kw + 2, npix = 18 * pitch;
    int cs = npix;
    // 8 threads of a store phase cover (chunk j, pixel p) pairs: make j*CS + p distinct mod 8
    const int want = nchunk >= 8 ? 1 : (nchunk == 4 ? 2 : (nchunk == 2 ? 4 : 0));
    while (want && (nchunk >= 8 ? (cs % 2 == 0) : (cs % 8 != want))) cs++;
    const size_t a_bytes = ((size_t)nchunk * cs * 16 + 127) & ~(size_t)127;
    if (2 * a_bytes <= budget) {
      P.KW = kw; P.pitch = pitch; P.npix = npix; P.CS = cs; P.a_bytes = (int)a_bytes;
      P.nstages = (int)std::min<size_t>(V2_MAX_STAGES, budget / a_bytes);
      break;
    }
  }
  if (!P.KW) VFI_FAIL(h, VFI_E_ARG, "conv_v2: no tile configuration fits (Cin=%d Cout=%d)", L.cin, L.cout);
  P.acc_stride = P.KW * L.cout;
  int cols = 2 * P.acc_stride;
  P.tmem_cols = cols <= 32 ? 32 : cols <= 64 ? 64 : cols <= 128 ? 128 : cols <= 256 ? 256 : 512;
  P.tiles_x = cdiv(W, 8 * P.KW);
  P.tiles_y = cdiv(H, 16);
  P.num_tiles = N * P.tiles_x * P.tiles_y;
  const size_t smem = (size_t)P.w_bytes + ((2 * L.cout * 4 + 127) & ~127) + (size_t)P.nstages * P.a_bytes;
  static bool attr_set = false;
  if (!attr_set) {
    VFI_CUDA(h, cudaFuncSetAttribute(conv_v2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024));
    attr_set = true;
  }
  const int grid = std::min(P.num_tiles, h->sm_count);
  conv_v2_kernel<<<grid, V2_THREADS, smem, s>>>(P);
  VFI_LAUNCHED(h);
  return 0;
}
