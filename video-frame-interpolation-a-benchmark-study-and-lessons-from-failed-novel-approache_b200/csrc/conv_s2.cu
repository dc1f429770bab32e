// conv_s2.cu -- resident-weight implicit-GEMM kernel for IFBlock.conv0's two stride-2 3x3 convolutions
// (conv(in_planes, c//2, 3, 2, 1), conv(c//2, c, 3, 2, 1), each + LeakyReLU(0.2)).
//
// The input arrives as the four stride-2 phase planes [N, 4, H/2, W/2, Cin] its producer wrote (plane = 2*(y&1) +
// (x&1)), so every tap of a stride-2 conv is a unit-stride window of ONE plane.  Per macro tile of 16 x 8KW output
// pixels the producer warps copy the 17 x (8KW+1) window of each plane into shared memory ONCE (cp.async 16 B, zero
// fill = conv padding), laid out [Cin/8][plane][17][8KW+1][8 ch] -- the canonical no-swizzle K-major UMMA layout whose
// 8-row group stride (SBO) is the plane row pitch -- and all nine taps are tcgen05.mma issued from that buffer with
// shifted start addresses: tap (ky,kx) -> plane (ky!=1, kx!=1), window offset (ky!=0, kx!=0).  Weights stay resident
// in shared memory for the whole persistent CTA; accumulators are double buffered in TMEM.
// Warp roles (544 threads): warps 0-7 cp.async producers (their copy lists -- smem offset, source offset, window
// coordinates -- are tile independent and live in registers), warp 8 MMA issuer, warps 9-16 epilogue.
#include <algorithm>

#include "umma.cuh"

using namespace umma;

struct S2Params {
  const bf16* in;   // [N, 4, Ho, Wo, CIN] phase planes (Ho = Hin/2, Wo = Win/2)
  bf16* out;        // [N, Ho, Wo, COUT] or, out_phase, [N, 4, Ho/2, Wo/2, COUT]
  const bf16* wpack;
  const float* bias;
  int N, Ho, Wo;
  int tiles_x, tiles_y, num_tiles;
  int nstages, out_phase, act;
  float slope;
};

#define S2_PRODUCERS 256
#define S2_THREADS (S2_PRODUCERS + 32 + 256)
#define S2_PW (S2_PRODUCERS / 32)  // producer warps; warp S2_PW = MMA issuer; the 8 warps after it = epilogue
#define S2_MAX_STAGES 6

template <int CIN, int COUT, int KW>
struct S2Cfg {
  static constexpr int NCHUNK = CIN / 8;
  static constexpr int PW = 8 * KW + 1;  // plane window pitch (pixels)
  static constexpr int PH = 17;
  static constexpr int PLANE_PIX = PH * PW;
  static constexpr int NPIX = 4 * PLANE_PIX;
  static constexpr int CS = chunk_stride(NPIX, NCHUNK);
  static constexpr int A_BYTES = (NCHUNK * CS * 16 + 127) & ~127;
  static constexpr int W_BYTES = 9 * CIN * COUT * 2;
  static constexpr int B_BYTES = (COUT * 4 + 127) & ~127;
  static constexpr int ACC_STRIDE = KW * COUT;
  static constexpr int TMEM_COLS = 2 * ACC_STRIDE <= 32 ? 32 : 2 * ACC_STRIDE <= 64 ? 64 : 2 * ACC_STRIDE <= 128 ? 128
                                   : 2 * ACC_STRIDE <= 256 ? 256 : 512;
  static constexpr int TOTAL_ITEMS = NPIX * NCHUNK;
  static_assert(2 * ACC_STRIDE <= 512, "TMEM budget");
  static_assert(COUT % 16 == 0 && CIN % 16 == 0, "MMA shape");
};

template <int CIN, int COUT, int KW>
__global__ void __launch_bounds__(S2_THREADS, 1) conv_s2_kernel(const __grid_constant__ S2Params P) {
  using C = S2Cfg<CIN, COUT, KW>;
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint64_t full_bar[S2_MAX_STAGES], empty_bar[S2_MAX_STAGES], tfull_bar[2], tempty_bar[2];
  __shared__ uint32_t tmem_base_smem;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint8_t* w_s = smem;
  float* bias_s = reinterpret_cast<float*>(smem + C::W_BYTES);
  uint8_t* a_s = smem + C::W_BYTES + C::B_BYTES;

  if (threadIdx.x == 0) {
    for (int i = 0; i < P.nstages; i++) { bar_init(&full_bar[i], S2_PRODUCERS / 32); bar_init(&empty_bar[i], 1); }
    for (int i = 0; i < 2; i++) { bar_init(&tfull_bar[i], 1); bar_init(&tempty_bar[i], 8); }
    fence_barrier_init();
  }
  if (warp == S2_PW) tmem_alloc(&tmem_base_smem, C::TMEM_COLS);
  {
    const uint4* src = reinterpret_cast<const uint4*>(P.wpack);
    uint4* dst = reinterpret_cast<uint4*>(w_s);
    for (int i = threadIdx.x; i < C::W_BYTES / 16; i += S2_THREADS) dst[i] = __ldg(src + i);
    for (int i = threadIdx.x; i < COUT; i += S2_THREADS) bias_s[i] = P.bias[i];
    fence_async_smem();
  }
  tc_before();
  __syncthreads();
  tc_after();
  const uint32_t tmem_base = tmem_base_smem;

  if (warp < S2_PW) {
    // ================= producers: 4 plane windows -> smem [chunk][plane][17][PW][8ch] =================
    constexpr int PER_THREAD = (C::TOTAL_ITEMS + S2_PRODUCERS - 1) / S2_PRODUCERS;
    // The copy list of a thread does not depend on the tile: item u = (chunk j, plane, window row pr, window col pc).
    //   meta = smem offset (16-byte units, bits 0-13) | pr << 14 | pc << 19 | skip << 31;  srel = source element offset
    //   relative to the window origin of plane 0.
    uint32_t meta[PER_THREAD];
    int srel[PER_THREAD];
#pragma unroll
    for (int u = 0; u < PER_THREAD; u++) {
      const int i = (int)threadIdx.x + u * S2_PRODUCERS;
      const int pix = i / C::NCHUNK, j = i - pix * C::NCHUNK;
      const int plane = pix / C::PLANE_PIX, rem = pix - plane * C::PLANE_PIX;
      const int pr = rem / C::PW, pc = rem - pr * C::PW;
      // even-row planes never use window row 0, even-column planes never use window column 0
      const bool skip = i >= C::TOTAL_ITEMS || (pr == 0 && plane < 2) || (pc == 0 && !(plane & 1));
      meta[u] = (uint32_t)(j * C::CS + pix) | ((uint32_t)pr << 14) | ((uint32_t)pc << 19) | (skip ? 0x80000000u : 0u);
      srel[u] = ((plane * P.Ho + pr) * P.Wo + pc) * CIN + j * 8;
    }
    // tiles in flight per thread; nstages - depth - 1 tiles of slack keep the publish hand-shake off the MMA warp's
    // critical path (see conv_v2.cu)
    const int depth = P.nstages - 2 < 1 ? 1 : (P.nstages - 2 > 3 ? 3 : P.nstages - 2);
    int it = 0;
    TileWalk tw((int)blockIdx.x, (int)gridDim.x, P.tiles_x, P.tiles_y);   // no per-tile integer divisions (umma.cuh)
    StageWalk sw(P.nstages), pub(P.nstages);                               // pub: the stage published `depth` iterations later
    for (; tw.tile < P.num_tiles; tw.next(), sw.next(), it++) {
      const int st = sw.st;
      const uint32_t ph = sw.ph;
      const int tx = tw.tx, ty = tw.ty, n = tw.n;
      const int c0 = tx * 8 * KW - 1, r0 = ty * 16 - 1;  // plane coordinates of the window origin
      if (it >= depth) {
        if (depth == 3) cp_wait<2>(); else if (depth == 2) cp_wait<1>(); else cp_wait<0>();
        fence_async_smem();
        __syncwarp();
        if (lane == 0) bar_arrive(&full_bar[pub.st]);
        pub.next();
      }
      bar_wait(&empty_bar[st], ph ^ 1);
      const uint32_t dst = s_u32(a_s + (size_t)st * C::A_BYTES);
      const bf16* img = P.in + (size_t)n * 4 * P.Ho * P.Wo * CIN;
      const bf16* org = img + ((ptrdiff_t)r0 * P.Wo + c0) * CIN;  // window origin in plane 0 (may lie outside the image)
#pragma unroll
      for (int u = 0; u < PER_THREAD; u++) {
        const uint32_t mt = meta[u];
        if (!(mt & 0x80000000u)) {
          const int R = r0 + (int)((mt >> 14) & 31), Cc = c0 + (int)((mt >> 19) & 63);
          const bool ok = (unsigned)R < (unsigned)P.Ho && (unsigned)Cc < (unsigned)P.Wo;
          cp_async16(dst + ((mt & 0x3FFFu) << 4), ok ? org + srel[u] : img, ok ? 16u : 0u);
        }
      }
      cp_commit();
    }
    cp_wait<0>();
    fence_async_smem();
    __syncwarp();
    if (lane == 0)
      for (int k = (it < depth ? it : depth); k > 0; k--) bar_arrive(&full_bar[(it - k) % P.nstages]);
  } else if (warp == S2_PW) {
    // ================= MMA issuer =================
    const uint32_t leader = elect_one();
    constexpr uint32_t IDESC = idesc_bf16_m128(COUT);
    constexpr uint32_t A_LBO = C::CS * 16, A_SBO = C::PW * 16, B_LBO = COUT * 16, B_SBO = 128;
    constexpr int KSTEPS = CIN / 16;
    const uint64_t bd0 = desc_ns(s_u32(w_s), B_LBO, B_SBO);
    int it = 0;
    StageWalk sw(P.nstages);
    for (int tile = blockIdx.x; tile < P.num_tiles; tile += gridDim.x, sw.next(), it++) {
      const int st = sw.st;
      const uint32_t ph = sw.ph;
      const int as = it & 1;
      bar_wait(&tempty_bar[as], (uint32_t)(((it >> 1) & 1) ^ 1));
      bar_wait(&full_bar[st], ph);
      tc_after();
      const uint64_t ad0 = desc_ns(s_u32(a_s + (size_t)st * C::A_BYTES), A_LBO, A_SBO);
      const uint32_t dcol0 = tmem_base + (uint32_t)(as * C::ACC_STRIDE);
#pragma unroll
      for (int tap = 0; tap < 9; tap++) {
#pragma unroll
        for (int k = 0; k < KSTEPS; k++) {
#pragma unroll
          for (int j = 0; j < KW; j++) {
            const int ky = tap / 3, kx = tap % 3;
            const int plane = (ky != 1 ? 2 : 0) + (kx != 1 ? 1 : 0);
            const uint32_t a_off =
                (uint32_t)(plane * C::PLANE_PIX + (ky != 0 ? 1 : 0) * C::PW + (kx != 0 ? 1 : 0) + 8 * j + 2 * k * C::CS);
            const uint32_t b_off = (uint32_t)((tap * C::NCHUNK + 2 * k) * COUT);
            mma_bf16_elect(leader, dcol0 + (uint32_t)(j * COUT), ad0 + a_off, bd0 + b_off, IDESC, (tap | k) ? 1u : 0u);
          }
        }
      }
      if (leader) {
        commit_to(&empty_bar[st]);
        commit_to(&tfull_bar[as]);
      }
      __syncwarp();
    }
  } else {
    // ================= epilogue: 8 warps, two per TMEM lane group =================
    const int lg = warp & 3;
    const int half = (warp - S2_PW - 1) >> 2;
    const int m = lg * 32 + lane;
    const int ly = m >> 3, lx = m & 7;
    const float4* bias4 = reinterpret_cast<const float4*>(bias_s);
    int it = 0;
    TileWalk tw((int)blockIdx.x, (int)gridDim.x, P.tiles_x, P.tiles_y);
    for (; tw.tile < P.num_tiles; tw.next(), it++) {
      const int as = it & 1;
      const int tx = tw.tx, ty = tw.ty, n = tw.n;
      const int y = ty * 16 + ly;
      bar_wait(&tfull_bar[as], (uint32_t)((it >> 1) & 1));
      tc_after();
#pragma unroll
      for (int jj = 0; jj < (KW + 1) / 2; jj++) {
        const int j = 2 * jj + half;
        if (j < KW) {
          const int x = (tx * KW + j) * 8 + lx;
          const bool valid = (y < P.Ho) && (x < P.Wo);
          const uint32_t taddr = tmem_base + ((uint32_t)(lg * 32) << 16) + (uint32_t)(as * C::ACC_STRIDE + j * COUT);
          size_t opix;
          if (!P.out_phase) opix = ((size_t)n * P.Ho + y) * P.Wo + x;
          else opix = ((((size_t)n * 4 + (y & 1) * 2 + (x & 1)) * (P.Ho >> 1) + (y >> 1)) * (P.Wo >> 1) + (x >> 1));
          bf16* op = P.out + opix * COUT;
#pragma unroll
          for (int c = 0; c < COUT; c += 16) {
            float v[16];
            ld16(taddr + c, v);
            if (valid) {
#pragma unroll
              for (int q = 0; q < 4; q++) {
                const float4 b = bias4[(c >> 2) + q];
                v[4 * q] += b.x; v[4 * q + 1] += b.y; v[4 * q + 2] += b.z; v[4 * q + 3] += b.w;
              }
              if (P.act) {
#pragma unroll
                for (int i = 0; i < 16; i++) v[i] = v[i] > 0.f ? v[i] : v[i] * P.slope;
              }
              uint4 o0, o1;
              o0.x = pk_bf16(v[0], v[1]); o0.y = pk_bf16(v[2], v[3]); o0.z = pk_bf16(v[4], v[5]); o0.w = pk_bf16(v[6], v[7]);
              o1.x = pk_bf16(v[8], v[9]); o1.y = pk_bf16(v[10], v[11]); o1.z = pk_bf16(v[12], v[13]); o1.w = pk_bf16(v[14], v[15]);
              st_v8(op + c, o0, o1);
            }
          }
        }
      }
      tc_before();
      __syncwarp();
      if (lane == 0) bar_arrive(&tempty_bar[as]);
    }
  }
  tc_before();
  __syncthreads();
  if (warp == S2_PW) {
    tc_after();
    tmem_dealloc(tmem_base, C::TMEM_COLS);
  }
}

// =============================================================================================
bool s2_supported(const Layer& L) {
  if (L.kind != LK_CONV3 || L.stride != 2 || L.cin_pad < 16) return false;
  return (L.cin_pad == 32 && (L.cout == 16 || L.cout == 32 || L.cout == 48 || L.cout == 64)) ||
         (L.cin_pad == 16 && L.cout == 32) || (L.cin_pad == 48 && L.cout == 96);
}

// Conv2d weights [Co][Ci][3][3] -> resident smem image [tap][Cin_pad/8][Co][8] bf16 (padded input channels are zero)
int s2_build_layer(vfi_handle* h, Layer& L, const std::vector<float>& w) {
  if (!s2_supported(L)) return 0;
  const int nch = L.cin_pad / 8;
  std::vector<uint16_t> pk((size_t)9 * L.cin_pad * L.cout, 0);
  for (int tap = 0; tap < 9; tap++)
    for (int ci = 0; ci < L.cin; ci++)
      for (int co = 0; co < L.cout; co++)
        pk[(((size_t)tap * nch + ci / 8) * L.cout + co) * 8 + ci % 8] = f2bf_host(w[((size_t)co * L.cin + ci) * 9 + tap]);
  VFI_CUDA(h, cudaMalloc(&L.ws2, pk.size() * 2));
  VFI_CUDA(h, cudaMemcpy(L.ws2, pk.data(), pk.size() * 2, cudaMemcpyHostToDevice));
  return 0;
}

template <int CIN, int COUT, int KW>
static int launch_s2(vfi_handle* h, S2Params& P, cudaStream_t s) {
  using C = S2Cfg<CIN, COUT, KW>;
  const size_t fixed = (size_t)C::W_BYTES + C::B_BYTES;
  const size_t budget = 220 * 1024 - fixed;
  P.nstages = (int)std::min<size_t>(S2_MAX_STAGES, budget / C::A_BYTES);
  if (P.nstages < 2) VFI_FAIL(h, VFI_E_ARG, "conv_s2: shared memory budget (Cin=%d)", CIN);
  P.tiles_x = cdiv(P.Wo, 8 * KW);
  P.tiles_y = cdiv(P.Ho, 16);
  P.num_tiles = P.N * P.tiles_x * P.tiles_y;
  const size_t smem = fixed + (size_t)P.nstages * C::A_BYTES;
  VFI_CUDA(h, cudaFuncSetAttribute(conv_s2_kernel<CIN, COUT, KW>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   224 * 1024));
  const int grid = std::min(P.num_tiles, h->sm_count);
  conv_s2_kernel<CIN, COUT, KW><<<grid, S2_THREADS, smem, s>>>(P);
  VFI_LAUNCHED(h);
  return 0;
}

// in: phase planes of the [N, Hin, Win, cin_pad] input; out: [N, Hin/2, Win/2, cout] plain or phase planes
int k_conv_s2(vfi_handle* h, const Layer& L, const bf16* in, bf16* out, int N, int Hin, int Win, bool out_phase,
              float slope, cudaStream_t s) {
  ProfScope _ps(h, PC_CONV_TC, s);
  if (!L.ws2) VFI_FAIL(h, VFI_E_STATE, "conv_s2: layer not packed");
  if ((Hin | Win) & 1) VFI_FAIL(h, VFI_E_ARG, "stride-2 conv needs even input dims");
  S2Params P;
  memset(&P, 0, sizeof(P));
  P.in = in; P.out = out; P.wpack = L.ws2; P.bias = L.bias;
  P.N = N; P.Ho = Hin / 2; P.Wo = Win / 2;
  P.out_phase = out_phase ? 1 : 0;
  if (out_phase && ((P.Ho | P.Wo) & 1)) VFI_FAIL(h, VFI_E_ARG, "phase-plane output needs even dims");
  P.act = L.act ? 1 : 0;
  P.slope = slope;
  // KW chosen so that 4 stages (3 tiles of cp.async in flight per producer thread) fit: these layers are HBM-bound
  if (L.cin_pad == 32 && L.cout == 16) return launch_s2<32, 16, 1>(h, P, s);
  if (L.cin_pad == 32 && L.cout == 32) return launch_s2<32, 32, 1>(h, P, s);
  if (L.cin_pad == 32 && L.cout == 48) return launch_s2<32, 48, 1>(h, P, s);
  if (L.cin_pad == 32 && L.cout == 64) return launch_s2<32, 64, 1>(h, P, s);
  if (L.cin_pad == 48 && L.cout == 96) return launch_s2<48, 96, 1>(h, P, s);
  if (L.cin_pad == 16 && L.cout == 32) return launch_s2<16, 32, 2>(h, P, s);
  VFI_FAIL(h, VFI_E_ARG, "conv_s2: no instance for Cin=%d Cout=%d", L.cin_pad, L.cout);
}
