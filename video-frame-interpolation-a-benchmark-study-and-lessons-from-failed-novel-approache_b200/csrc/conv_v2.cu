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
// Warp roles (544 threads): warps 0-7 producers (cp.async 16 B, zero fill = conv padding, up to 3 tiles in flight; the
// per-thread copy list is tile independent and lives in registers), warp 8 TMEM allocator + MMA issuer (converged
// warp, elected lane), warps 9-16 epilogue (two per TMEM lane group).
// Layer semantics: SURVEY.md Appendix A.2 (conv(), ResConv).
#include <algorithm>

#include <cuda.h>
#include "common.cuh"
#include "umma.cuh"
#include <cuda_fp16.h>

struct V2Params {
  const bf16* in;
  bf16* out;
  const bf16* wpack;
  const float* bias;
  const float* beta;  // null => plain conv
  int N, H, W;
  int tiles_x, tiles_y, num_tiles;
  int nstages;
  int res_global;  // ResConv residual from global memory (L2) instead of the smem halo tile (measured 2.5 % slower)
  int ablate;  // debug (vfi_set_option "v2_ablate"): 1 skip loads, 2 skip MMAs, 4 skip epilogue math+stores, 8 skip TMEM loads
  int act;
  float slope;
  // bias / beta by output column, read by the epilogue through the constant bank: the shared-memory pipe is saturated
  // by the tensor core's operand reads while MMAs are in flight
  float bias_c[128];
  float beta_c[128];
  float* low_out;  // EPI 2 (final block's lastconv): [N, 4H, 4W, low_c] fp32, channels 0-4 written
  int low_c;
  unsigned long long* trace;  // debug (vfi_set_option "v2_trace"): CTA 0 timestamps [role][tile < 64][8]
};

#define V2_PRODUCERS 256
#define V2_EPI_WARPS 16
#define V2_THREADS (V2_PRODUCERS + 32 + 32 * V2_EPI_WARPS)
#define V2_PW (V2_PRODUCERS / 32)  // producer warps; warp V2_PW = MMA issuer; the V2_EPI_WARPS warps after it = epilogue
#define V2_MAX_STAGES 6
#define V2_DEPTH 3
#ifndef V2_NACC_MAX
#define V2_NACC_MAX 4
#endif
#ifndef V2_TMA_CHUNKS
// chunks (of NCHUNK = Cin/8) a TMA instance loads by TMA; the rest would come through the cp.async producers.  All of
// them: splitting the tile half / half between the two paths measured 8-12 % SLOWER than TMA alone on every instance.
#define V2_TMA_CHUNKS(n) (n)
#endif
#ifndef V2_PM_BASEOFF
#define V2_PM_BASEOFF 0
#endif
#ifndef V2_EPI_GROUPS
// 1: all 16 epilogue warps work on one tile, then all move to the next (one tile generation in flight).
// 2: two groups of 8 warps take ALTERNATE tiles, each warp doing in two sequential halves what two warps do in mode 1: two
//    generations in flight hide the epilogue's serial chain (wait -> residual LDS -> tcgen05.ld -> math -> stores), which
//    tools/v2_trace.py shows to be the tile period of the c = 32 / 64 layers.  Not used by the final lastconv (EPI 2).
#define V2_EPI_GROUPS 2
#endif
#ifndef V2_TMA_SLIM
#define V2_TMA_SLIM 0  // 1: TMA instances run 576 threads (one producer warp) with bias / beta in registers (measured 2 % slower)
#endif

// Compile-time geometry of one (Cin, Cout, KW) instance.
__device__ __forceinline__ unsigned long long v2_now() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%clock64;" : "=l"(t));  // SM cycle counter: ~20 cycles a stamp (%globaltimer cost 0.3-0.5 us)
  return t;
}
#define V2_TRACE(role, tile_i, ev)                                                              \
  do {                                                                                          \
    if (P.trace && blockIdx.x == 0 && (tile_i) < 64) P.trace[((role)*64 + (tile_i)) * 8 + (ev)] = v2_now(); \
  } while (0)

// COUT here = output channels computed by ONE CTA (the layer's Cout / NSPLIT)
template <int CIN, int COUT, int KW, bool TMA = false, bool PM = false>
struct V2Cfg {
  // PM (pixel-major): the halo tile is [Cin/64][pixel][64 ch], one 128-byte row per pixel and 64-channel block, written by
  // ONE TMA box per block with the 128-byte swizzle (128-byte box rows instead of 16-byte ones), read by K-major SWIZZLE_128B
  // descriptors whose start address is shifted per tap (base_offset keeps the swizzle phase)
  static constexpr int NKB = CIN / 64;
  static constexpr int NCHUNK = CIN / 8;
  static constexpr int PITCH = 8 * KW + 2;    // halo row pitch in pixels
  static constexpr int NPIX = 18 * PITCH;
  // chunk stride in 16-byte units: >= NPIX, padded so the 8 lanes of a 16-byte store phase (chunk j, pixel p) hit
  // 8 distinct bank groups: (j*CS + p) mod 8 distinct
  // (TMA loads: every chunk is one tensor box whose destination must be 128-byte aligned -> a multiple of 8)
  static constexpr int cs() {
    int c = NPIX;
    if (TMA) { while (c % 8) c++; return c; }
    if (NCHUNK >= 8) { while (c % 2 == 0) c++; }
    else if (NCHUNK == 4) { while (c % 8 != 2) c++; }
    else if (NCHUNK == 2) { while (c % 8 != 4) c++; }
    return c;
  }
  static constexpr int CS = cs();
  static constexpr int A_BYTES = PM ? ((NKB * NPIX * 128 + 1023) & ~1023) : ((NCHUNK * CS * 16 + 127) & ~127);
  static constexpr int W_BYTES = 9 * CIN * COUT * 2;
  static constexpr int BB_BYTES = PM ? 1024 : ((2 * COUT * 4 + 127) & ~127);
  static_assert(!PM || (TMA && CIN % 64 == 0 && W_BYTES % 1024 == 0), "pixel-major tiles: 64-channel blocks, 1024-byte aligned stages");
  static constexpr int ACC_COLS = KW * COUT;  // accumulator columns of one tile
  // TMEM accumulator ring: up to four buffers, each on a power-of-two column stride.  Two were not enough: handing a
  // buffer from the MMA warp to the epilogue and back (commit -> mbarrier -> tcgen05.ld -> arrive -> poll) takes ~2.9 us
  // with no work at all (v2_ablate = 7: 1.45 us per tile), longer than the tile's 1.7 us of MMAs
  static constexpr int ACC_STRIDE = ACC_COLS <= 32 ? 32 : ACC_COLS <= 64 ? 64 : ACC_COLS <= 128 ? 128 : 256;
  static constexpr int NACC = 512 / ACC_STRIDE < V2_NACC_MAX ? 512 / ACC_STRIDE : V2_NACC_MAX;
  static constexpr int TMEM_COLS = NACC * ACC_STRIDE < 32 ? 32 : NACC * ACC_STRIDE;
  static constexpr int TOTAL_ITEMS = NPIX * NCHUNK;
  static constexpr int UNROLL = (TOTAL_ITEMS + 2 * V2_PRODUCERS - 1) / (2 * V2_PRODUCERS);  // two load batches per tile
  static_assert(2 * ACC_COLS <= 512, "TMEM budget");
};

namespace {
using namespace umma;  // mbarrier / cp.async / tcgen05 wrappers and the no-swizzle descriptor shared with conv_s2.cu and the fused gathers
// polling wait with back-off, for waits that are normally long (stage release, accumulator release): a warp parked here
// must not compete with the epilogue warps for issue slots and the shared-memory pipe
__device__ __forceinline__ void bar_wait_relaxed(uint64_t* b, uint32_t parity) {
  uint32_t done;
  do {
    asm volatile(
        "{\n\t.reg .pred P1;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, P1;\n\t}"
        : "=r"(done)
        : "r"(s_u32(b)), "r"(parity)
        : "memory");
    if (!done) __nanosleep(200);
  } while (!done);
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
// tcgen05.ld of N consecutive fp32 accumulator columns of this warp's 32 lanes; NO wait (caller issues wait::ld)
template <int N>
__device__ __forceinline__ void ld_cols(uint32_t taddr, float* v) {
  static_assert(N == 8 || N == 16 || N == 32, "ld_cols width");
  uint32_t* r = reinterpret_cast<uint32_t*>(v);
  if constexpr (N == 8) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(taddr)
                 : "memory");
  } else if constexpr (N == 16) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
  } else {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
  }
}
// one 32-byte (full-sector) global store
__device__ __forceinline__ void st_global_v8(float* p, float a, float b, float c, float d, float e, float f, float g,
                                             float h) {
  asm volatile("st.global.v8.f32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d),
               "f"(e), "f"(f), "f"(g), "f"(h)
               : "memory");
}
__device__ __forceinline__ uint32_t idesc_bf16(int n) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | (8u << 24);  // F32 acc, BF16 x BF16, M = 128
}
}  // namespace

// EPI 0: conv (+ResConv) -> bf16 NHWC.  EPI 1: Head.cnn3 = ConvTranspose2d(16,4,4,2,1) run as a 3x3 conv whose 16
// output columns are (output phase p,q) x (4 channels); the epilogue scatters them to the 2x2 output quad.
// NSPLIT > 1: the layer's output channels are split over NSPLIT groups of CTAs (CTA c serves group c % NSPLIT) so that
// the group's 9 x CIN x COUT/NSPLIT weight slice stays resident where the whole matrix would not fit (c = 96, 128).
// TMA: the halo tile arrives as NCHUNK tensor boxes {8 ch, PITCH, 18 rows} issued by ONE thread (zero fill outside the image
// = the padding) instead of ~10 cp.async per producer thread: the tile's 16-byte requests no longer queue in the LSU in
// front of the epilogue's loads and stores (tools/v2_trace.py: issuing them took 0.8-2.2 us of a 3 us tile period).
// TMA instances run one producer warp instead of eight: 576 threads, so the epilogue warps may use up to 112 registers and
// keep their bias / beta slice in registers (an indexed constant-bank load per use cost ~0.5 us per tile, v2_trace).
template <int CIN, int COUT, int KW, int EPI, int NSPLIT = 1, bool TMA = false, bool PM = false>
__global__ void __launch_bounds__((TMA && V2_TMA_SLIM) ? 32 + 32 + 32 * V2_EPI_WARPS : V2_THREADS, 1)
conv_v2_kernel(const __grid_constant__ V2Params P, const __grid_constant__ CUtensorMap tmA) {
  constexpr int PW_ = (TMA && V2_TMA_SLIM) ? 1 : V2_PW;                              // producer warps; warp PW_ issues the MMAs
  constexpr int NTHREADS = (TMA && V2_TMA_SLIM) ? 32 + 32 + 32 * V2_EPI_WARPS : V2_THREADS;
  constexpr int CN = COUT / NSPLIT;  // GEMM N of this CTA
  static_assert(COUT % NSPLIT == 0 && CN % 16 == 0, "channel split");
  using C = V2Cfg<CIN, CN, KW, TMA, PM>;
  constexpr int TJ = TMA ? V2_TMA_CHUNKS(C::NCHUNK) : 0;  // chunks loaded by TMA
  constexpr bool ALL_TMA = TMA && TJ == C::NCHUNK;
  static_assert(!(V2_TMA_SLIM && TMA) || ALL_TMA, "the slim variant has no cp.async producers");
  const int split = NSPLIT > 1 ? (int)blockIdx.x % NSPLIT : 0;
  const int co0 = split * CN;  // first output channel of this CTA
  const int tile0 = (int)blockIdx.x / NSPLIT, tile_step = (int)gridDim.x / NSPLIT;
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t full_bar[V2_MAX_STAGES], empty_bar[V2_MAX_STAGES], tfull_bar[4], tempty_bar[4];
  __shared__ uint32_t tmem_base_smem;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint8_t* w_s = smem;
  float* bias_s = reinterpret_cast<float*>(smem + C::W_BYTES);
  float* beta_s = bias_s + CN;
  uint8_t* a_s = smem + C::W_BYTES + C::BB_BYTES;
  const bool use_res = EPI == 0 && P.beta != nullptr;

  // ---- one-time setup: barriers, TMEM, resident weights ----
  if (threadIdx.x == 0) {
    // epilogue warps that arrive per tile: all 16, or the 8 of the group that owns the tile
    constexpr int EPI_ARRIVE = (V2_EPI_GROUPS == 2 && EPI != 2) ? V2_EPI_WARPS / 2 : V2_EPI_WARPS;
    for (int i = 0; i < P.nstages; i++) { bar_init(&full_bar[i], ALL_TMA ? 1 : V2_PRODUCERS / 32 + (TJ > 0 ? 1 : 0)); bar_init(&empty_bar[i], (use_res && !P.res_global) ? 1 + EPI_ARRIVE : 1); }
    for (int i = 0; i < C::NACC; i++) { bar_init(&tfull_bar[i], 1); bar_init(&tempty_bar[i], EPI_ARRIVE); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == PW_) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s_u32(&tmem_base_smem)),
                 "r"((uint32_t)C::TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  {
    const uint4* src = reinterpret_cast<const uint4*>(P.wpack + (size_t)split * 9 * CIN * CN);
    uint4* dst = reinterpret_cast<uint4*>(w_s);
    for (int i = threadIdx.x; i < C::W_BYTES / 16; i += NTHREADS) dst[i] = __ldg(src + i);
    for (int i = threadIdx.x; i < CN; i += NTHREADS) {
      bias_s[i] = P.bias[co0 + i];
      beta_s[i] = use_res ? P.beta[co0 + i] : 1.f;
    }
    fence_async_smem();
  }
  tc_before();
  __syncthreads();
  tc_after();
  const uint32_t tmem_base = tmem_base_smem;

  if (ALL_TMA && warp < PW_) {
    // ================= producer (TMA): one thread, NCHUNK tensor boxes per tile =================
    if (threadIdx.x == 0) {
      int it = 0;
      TileWalk tw(tile0, tile_step, P.tiles_x, P.tiles_y);
      StageWalk sw(P.nstages);
      for (; tw.tile < P.num_tiles; tw.next(), sw.next(), it++) {
        const int st = sw.st;
        const uint32_t ph = sw.ph;
        const int tx = tw.tx, ty = tw.ty, n = tw.n;
        V2_TRACE(0, it, 0);  // wants stage
        if (P.ablate & 256) bar_wait(&empty_bar[st], ph ^ 1); else bar_wait_relaxed(&empty_bar[st], ph ^ 1);
        V2_TRACE(0, it, 1);  // got stage
        const uint32_t dst = s_u32(a_s + (size_t)st * C::A_BYTES);
        const uint32_t bar = s_u32(&full_bar[st]);
        if (P.ablate & 1) {  // debug: no loads, the stage keeps whatever it held
          asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
          continue;
        }
        if constexpr (PM) {
          asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"((uint32_t)(C::NKB * C::NPIX * 128))
                       : "memory");
#pragma unroll
          for (int kb = 0; kb < C::NKB; kb++)
            asm volatile(
                "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(
                    dst + (uint32_t)(kb * C::NPIX * 128)),
                "l"((uint64_t)&tmA), "r"(bar), "r"(kb * 64), "r"(tx * 8 * KW - 1), "r"(ty * 16 - 1), "r"(n)
                : "memory");
        } else {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"((uint32_t)(C::NCHUNK * C::NPIX * 16))
                     : "memory");
#pragma unroll
        for (int j = 0; j < C::NCHUNK; j++)
          asm volatile(
              "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(
                  dst + (uint32_t)(j * C::CS * 16)),
              "l"((uint64_t)&tmA), "r"(bar), "r"(j * 8), "r"(tx * 8 * KW - 1), "r"(ty * 16 - 1), "r"(n)
              : "memory");
        }
        V2_TRACE(0, it, 2);  // loads issued
      }
    }
  } else if (warp < PW_) {
    // ================= producers: halo tile -> smem [chunk][pixel][8ch] =================
    // cp.async (LDGSTS, 16 B, L2-only) straight into the UMMA layout; zero fill (src-size 0) is the conv padding.
    // Up to V2_DEPTH tiles are in flight per thread: tile i is published (wait_group -> fence.proxy.async -> arrive)
    // while tiles i+1..i+V2_DEPTH are still streaming in.  The copy list of a thread does not depend on the tile:
    //   meta = smem offset (16-byte units, bits 0-13) | halo row << 14 | halo col << 19 | unused << 31
    //   srel = source element offset relative to the halo origin
    // chunks TJ .. NCHUNK-1 come through cp.async (all of them without TMA; the branch is dead for an all-TMA instance)
    constexpr int NC = C::NCHUNK - TJ > 0 ? C::NCHUNK - TJ : 1;
    constexpr int ROW_ITEMS = C::PITCH * NC;
    constexpr int TOTAL_ITEMS = C::NPIX * NC;
    constexpr int PER_THREAD = (TOTAL_ITEMS + V2_PRODUCERS - 1) / V2_PRODUCERS;
    uint32_t meta[PER_THREAD];
    int srel[PER_THREAD];
#pragma unroll
    for (int u = 0; u < PER_THREAD; u++) {
      const int i = (int)threadIdx.x + u * V2_PRODUCERS;
      const int py = i / ROW_ITEMS, rem = i - py * ROW_ITEMS;
      const int px = rem / NC, j = TJ + rem - px * NC;
      meta[u] = (uint32_t)(j * C::CS + py * C::PITCH + px) | ((uint32_t)py << 14) | ((uint32_t)px << 19) |
                (i >= TOTAL_ITEMS ? 0x80000000u : 0u);
      srel[u] = (py * P.W + px) * CIN + j * 8;
    }
    // tiles in flight per thread.  A tile is published when the producer comes round `depth` iterations later, i.e.
    // after it has been granted the stage of tile it+depth-nstages: nstages - depth - 1 tiles of slack keep that
    // hand-shake off the MMA warp's critical path (with no slack, load / MMA / epilogue times ADD instead of overlap).
    const int depth = P.nstages - 2 < 1 ? 1 : (P.nstages - 2 > V2_DEPTH ? V2_DEPTH : P.nstages - 2);
    int it = 0;
    TileWalk tw(tile0, tile_step, P.tiles_x, P.tiles_y);
    StageWalk sw(P.nstages);
    for (; tw.tile < P.num_tiles; tw.next(), sw.next(), it++) {
      const int st = sw.st;
      const uint32_t ph = sw.ph;
      const int tx = tw.tx, ty = tw.ty, n = tw.n;
      const int gx0 = tx * 8 * KW - 1, gy0 = ty * 16 - 1;
      if (it >= depth) {
        // publish the oldest in-flight tile BEFORE claiming a new stage, so publishing never waits on a stage release
        if (depth == 3) asm volatile("cp.async.wait_group 2;" ::: "memory");
        else if (depth == 2) asm volatile("cp.async.wait_group 1;" ::: "memory");
        else asm volatile("cp.async.wait_group 0;" ::: "memory");
        fence_async_smem();
        __syncwarp();
        if (lane == 0) bar_arrive(&full_bar[(it - depth) % P.nstages]);
        if (threadIdx.x == 0) V2_TRACE(0, it - depth, 3);  // published
      }
      if (threadIdx.x == 0) V2_TRACE(0, it, 0);  // wants stage
      bar_wait_relaxed(&empty_bar[st], ph ^ 1);
      if (threadIdx.x == 0) V2_TRACE(0, it, 1);  // got stage
      const uint32_t dst = s_u32(a_s + (size_t)st * C::A_BYTES);
      if (TJ > 0 && threadIdx.x == 0) {  // the tile's first TJ chunks: tensor boxes, counted on the same full barrier
        const uint32_t bar = s_u32(&full_bar[st]);
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"((uint32_t)(TJ * C::NPIX * 16)) : "memory");
#pragma unroll
        for (int j = 0; j < TJ; j++)
          asm volatile(
              "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(
                  dst + (uint32_t)(j * C::CS * 16)),
              "l"((uint64_t)&tmA), "r"(bar), "r"(j * 8), "r"(gx0), "r"(gy0), "r"(n)
              : "memory");
      }
      const bf16* img = P.in + (size_t)n * P.H * P.W * CIN;
      const bf16* org = img + ((ptrdiff_t)gy0 * P.W + gx0) * CIN;  // halo origin (may lie outside the image)
      if (!(P.ablate & 1)) {
#pragma unroll
        for (int u = 0; u < PER_THREAD; u++) {
          const uint32_t mt = meta[u];
          if (!(mt & 0x80000000u)) {
            const int gy = gy0 + (int)((mt >> 14) & 31), gx = gx0 + (int)((mt >> 19) & 63);
            const bool ok = (unsigned)gy < (unsigned)P.H && (unsigned)gx < (unsigned)P.W;
            cp_async16(dst + ((mt & 0x3FFFu) << 4), ok ? org + srel[u] : img, ok ? 16u : 0u);
          }
        }
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
      if (threadIdx.x == 0) V2_TRACE(0, it, 2);  // loads issued
    }
    // drain: publish the last `depth` tiles
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    fence_async_smem();
    __syncwarp();
    if (lane == 0)
      for (int k = (it < depth ? it : depth); k > 0; k--) bar_arrive(&full_bar[(it - k) % P.nstages]);
  } else if (warp == PW_) {
    // ================= MMA issuer: the warp stays converged, one elected lane issues =================
    const uint32_t leader = elect_one();
    constexpr uint32_t IDESC = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(CN >> 3) << 17) | (8u << 24);
    constexpr uint32_t A_LBO = C::CS * 16, A_SBO = C::PITCH * 16, B_LBO = CN * 16, B_SBO = 128;
    constexpr int KSTEPS = CIN / 16;
    const uint64_t bd0 = desc_ns(s_u32(w_s), B_LBO, B_SBO);
    int it = 0;
    StageWalk sw(P.nstages);
    for (int tile = tile0; tile < P.num_tiles; tile += tile_step, sw.next(), it++) {
      const int st = sw.st;
      const uint32_t ph = sw.ph;
      const int as = it % C::NACC;
      if (leader) V2_TRACE(1, it, 0);
      if (P.ablate & 512) bar_wait(&tempty_bar[as], (uint32_t)(((it / C::NACC) & 1) ^ 1)); else bar_wait_relaxed(&tempty_bar[as], (uint32_t)(((it / C::NACC) & 1) ^ 1));
      if (leader) V2_TRACE(1, it, 1);  // accumulator free
      bar_wait(&full_bar[st], ph);
      if (leader) V2_TRACE(1, it, 2);  // tile landed
      tc_after();
      const uint32_t a_base = s_u32(a_s + (size_t)st * C::A_BYTES);
      const uint64_t ad0 = desc_ns(a_base, A_LBO, A_SBO);
      if (!(P.ablate & 2)) {
      const uint32_t dcol0 = tmem_base + (uint32_t)(as * C::ACC_STRIDE);
      // consecutive MMAs target different accumulators (j innermost) so they pipeline; all descriptor offsets are
      // compile-time immediates (16-byte units added to the start-address field)
#pragma unroll
      for (int tap = 0; tap < 9; tap++) {
#pragma unroll
        for (int k = 0; k < KSTEPS; k++) {
#pragma unroll
          for (int j = 0; j < KW; j++) {
            constexpr int dummy = 0;
            (void)dummy;
            const uint32_t b_off = (uint32_t)((tap * C::NCHUNK + 2 * k) * CN);
            if constexpr (PM) {
              // K-major SWIZZLE_128B: rows = pixels (128 B apart), 8-row group = 8 consecutive x, next group = next y;
              // k-step = 32 B inside the row; base_offset = the row phase of the (unaligned) start inside the 1024 B pattern
              const uint32_t addr = a_base + (uint32_t)((k / 4) * C::NPIX * 128 + ((tap / 3) * C::PITCH + (tap % 3) + 8 * j) * 128 +
                                                         (k % 4) * 32);
              uint64_t d = (uint64_t)((addr & 0x3FFFF) >> 4) | ((uint64_t)1 << 16) | ((uint64_t)(C::PITCH * 8) << 32) |
                           ((uint64_t)1 << 46) | ((uint64_t)(V2_PM_BASEOFF ? ((addr >> 7) & 7) : 0) << 49) | ((uint64_t)2 << 61);
              mma_bf16_elect(leader, dcol0 + (uint32_t)(j * CN), d, bd0 + b_off, IDESC, (tap | k) ? 1u : 0u);
            } else {
            const uint32_t a_off = (uint32_t)((tap / 3) * C::PITCH + (tap % 3) + 8 * j + 2 * k * C::CS);
            mma_bf16_elect(leader, dcol0 + (uint32_t)(j * CN), ad0 + a_off, bd0 + b_off, IDESC, (tap | k) ? 1u : 0u);
            }
          }
        }
      }
      }
      if (leader) {
        if (P.ablate & 64) bar_arrive(&empty_bar[st]); else commit_to(&empty_bar[st]);
        if (P.ablate & 128) bar_arrive(&tfull_bar[as]); else commit_to(&tfull_bar[as]);
        V2_TRACE(1, it, 3);  // MMAs issued
      }
      __syncwarp();
    }
  } else if constexpr (EPI == 2) {
    // ================= epilogue of the final block's lastconv: ConvTranspose2d(32, 52, 4, 2, 1) + PixelShuffle(2), of
    // which only flow delta + mask (output channels 0-4 = deconv channels 0-19) are consumed.  Accumulator column =
    // (2p+q)*20 + (2i+jj)*5 + c: deconv phase (p,q), pixel-shuffle sub-position (i,jj), channel c -> output pixel
    // (4y + 2p + i, 4x + 2q + jj).  Warp role = (sub-tile j, phase row p): its 40 columns are the 2 x 4 output pixels of
    // rows 4y+2p and 4y+2p+1, columns 4x .. 4x+3: 128 contiguous bytes per row and thread.
    static_assert(EPI != 2 || (CN == 80 && KW == 2), "final lastconv instance");
    const int lg = warp & 3;
    const int role = (warp - PW_ - 1) >> 2;  // 0..3
    const int j = role & 1, pr = role >> 1;
    const int m = lg * 32 + lane;
    const int ly = m >> 3, lx = m & 7;
    const float* bias_r = P.bias_c + pr * 40;
    int it = 0;
    TileWalk tw(tile0, tile_step, P.tiles_x, P.tiles_y);
    for (; tw.tile < P.num_tiles; tw.next(), it++) {
      const int as = it % C::NACC;
      const int tx = tw.tx, ty = tw.ty, n = tw.n;
      const int y = ty * 16 + ly, x = (tx * KW + j) * 8 + lx;
      bar_wait(&tfull_bar[as], (uint32_t)((it / C::NACC) & 1));
      tc_after();
      float v[40];
      const uint32_t taddr = tmem_base + ((uint32_t)(lg * 32) << 16) + (uint32_t)(as * C::ACC_STRIDE + j * CN + pr * 40);
      ld_cols<32>(taddr, v);
      ld_cols<8>(taddr + 32, v + 32);
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      tc_before();
      __syncwarp();
      if (lane == 0) bar_arrive(&tempty_bar[as]);
      if (y < P.H && x < P.W) {
        const int H4 = 4 * P.H, W4 = 4 * P.W, lc = P.low_c;
#pragma unroll
        for (int i = 0; i < 2; i++) {
          float* row = P.low_out + (((size_t)n * H4 + 4 * y + 2 * pr + i) * W4 + 4 * x) * lc;
          if (lc == 4) {
            // fp16 texels {flow delta x4, mask, 0, 0, 0} (16 B): the row's four texels are two full sectors
#pragma unroll
            for (int q = 0; q < 2; q++) {
              uint4 t[2];
#pragma unroll
              for (int jj = 0; jj < 2; jj++) {
                const int c0 = q * 20 + (2 * i + jj) * 5;
                const __half2 a = __floats2half2_rn(v[c0] + bias_r[c0], v[c0 + 1] + bias_r[c0 + 1]);
                const __half2 b = __floats2half2_rn(v[c0 + 2] + bias_r[c0 + 2], v[c0 + 3] + bias_r[c0 + 3]);
                const __half2 c = __floats2half2_rn(v[c0 + 4] + bias_r[c0 + 4], 0.f);
                t[jj] = make_uint4(*reinterpret_cast<const uint32_t*>(&a), *reinterpret_cast<const uint32_t*>(&b),
                                   *reinterpret_cast<const uint32_t*>(&c), 0u);
              }
              st_v8(row + q * 8, t[0], t[1]);
            }
            continue;
          }
#pragma unroll
          for (int q = 0; q < 2; q++)
#pragma unroll
            for (int jj = 0; jj < 2; jj++) {
              const int c0 = q * 20 + (2 * i + jj) * 5;
              st_global_v8(row + (2 * q + jj) * lc, v[c0] + bias_r[c0], v[c0 + 1] + bias_r[c0 + 1], v[c0 + 2] + bias_r[c0 + 2],
                           v[c0 + 3] + bias_r[c0 + 3], v[c0 + 4] + bias_r[c0 + 4], 0.f, 0.f, 0.f);
            }
        }
      }
    }
  } else if constexpr (V2_EPI_GROUPS == 2) {
    // ================= epilogue, two tile generations in flight: group = (epilogue warp index) / 8 owns tiles it = group,
    // group + 2, ...; inside a group a warp (lane group lg, r2 = 0 / 1) does the work of mode 1's roles r2 and r2 + 2 as two
    // sequential halves, so its register footprint is mode 1's.  The smem stage goes back after the second half's residual
    // reads, the TMEM buffer after the second half's tcgen05.ld.
    const int lg = warp & 3;
    const int role4 = (warp - PW_ - 1) >> 2;     // 0..3
    const int grp = role4 >> 1, r2 = role4 & 1;
    constexpr int NSJ = KW >= 2 ? 2 : 1;
    constexpr int Q = 4 / NSJ;
    constexpr int CH = EPI == 1 ? 8 : (CN % Q == 0 && (CN / Q == 16 || CN / Q == 32)) ? CN / Q : 16;
    constexpr int NSC = CN / CH;
    static_assert(NSC <= Q && NSC * CH == CN, "epilogue slicing");
    constexpr int NJ = KW / NSJ;
    static_assert(CH == 8 || CH == 16 || CH == 32, "epilogue slice width");
    constexpr int NR = CH / 8;
    const int m = lg * 32 + lane;
    const int ly = m >> 3, lx = m & 7;
    int it = grp;
    TileWalk tw(tile0 + grp * tile_step, 2 * tile_step, P.tiles_x, P.tiles_y);
    StageWalk sw(P.nstages, grp);
    for (; tw.tile < P.num_tiles; tw.next(), sw.next(), sw.next(), it += 2) {
      const int st = sw.st;
      const int as = it % C::NACC;
      const uint32_t aph = (uint32_t)((it / C::NACC) & 1);
      const int tx = tw.tx, ty = tw.ty, n = tw.n;
      const int y = ty * 16 + ly;
      const bool smem_res = use_res && !P.res_global;
      if (smem_res) bar_wait(&full_bar[st], sw.ph);
      const uint4* a_st = reinterpret_cast<const uint4*>(a_s + (size_t)st * C::A_BYTES);
#pragma unroll
      for (int half = 0; half < 2; half++) {
        const int role = r2 + 2 * half;
        const int jc = role % NSJ;
        const int cb = (role / NSJ) * CH;
        const bool active = role / NSJ < NSC;
        const float* bias_r = P.bias_c + co0 + (active ? cb : 0);  // warp-uniform constant-bank reads
        const float* beta_r = P.beta_c + co0 + (active ? cb : 0);
        uint4 res[NJ * NR];
        if (use_res && P.res_global) {
          if (active)
#pragma unroll
          for (int jj = 0; jj < NJ; jj++) {
            const int x = (tx * KW + jj * NSJ + jc) * 8 + lx;
            const uint4* rp = reinterpret_cast<const uint4*>(P.in + (((size_t)n * P.H + y) * P.W + x) * CIN + co0 + cb);
            const bool ok = (y < P.H) && (x < P.W);
#pragma unroll
            for (int ch = 0; ch < NR; ch++) res[jj * NR + ch] = ok ? __ldg(rp + ch) : make_uint4(0u, 0u, 0u, 0u);
          }
        } else if (use_res) {
          if (active)
#pragma unroll
          for (int jj = 0; jj < NJ; jj++) {
            const int pc = (ly + 1) * C::PITCH + 8 * (jj * NSJ + jc) + lx + 1;  // this thread's centre pixel
#pragma unroll
            for (int ch = 0; ch < NR; ch++) {
              if constexpr (PM) {
                const int c8 = (co0 + cb) / 8 + ch;
                const uint32_t row = s_u32(a_s + (size_t)st * C::A_BYTES) + (uint32_t)((c8 >> 3) * C::NPIX * 128 + pc * 128);
                const uint32_t ad = row + (uint32_t)(((c8 & 7) ^ ((row >> 7) & 7)) << 4);
                asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                             : "=r"(res[jj * NR + ch].x), "=r"(res[jj * NR + ch].y), "=r"(res[jj * NR + ch].z), "=r"(res[jj * NR + ch].w)
                             : "r"(ad));
              } else {
                res[jj * NR + ch] = a_st[((co0 + cb) / 8 + ch) * C::CS + pc];
              }
            }
          }
          if (half == 1) {   // both halves' residuals are in registers: hand the stage back (fence: see mode 1)
            if (TMA) fence_async_smem();
            __syncwarp();
            if (lane == 0) bar_arrive(&empty_bar[st]);
          }
        }
        if (half == 0) { bar_wait(&tfull_bar[as], aph); tc_after(); }
        float v[NJ][CH];
        if (active) {
#pragma unroll
          for (int jj = 0; jj < NJ; jj++)
            ld_cols<CH>(tmem_base + ((uint32_t)(lg * 32) << 16) + (uint32_t)(as * C::ACC_STRIDE + (jj * NSJ + jc) * CN + cb), v[jj]);
          asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        }
        if (half == 1) {     // the buffer's last read: back to the MMA warp
          tc_before();
          __syncwarp();
          if (lane == 0) bar_arrive(&tempty_bar[as]);
        }
        if (active) {
#pragma unroll
          for (int jj = 0; jj < NJ; jj++) {
            const int j = jj * NSJ + jc;
            const int x = (tx * KW + j) * 8 + lx;
            if ((y < P.H) && (x < P.W)) {
#pragma unroll
              for (int i = 0; i < CH; i++) v[jj][i] += bias_r[i];
              if constexpr (EPI == 1) {
                static_assert(EPI != 1 || (COUT == 16 && CH == 8), "Head deconv instance");
                uint4 o;
                o.x = pk_bf16(v[jj][0], v[jj][1]); o.y = pk_bf16(v[jj][2], v[jj][3]);
                o.z = pk_bf16(v[jj][4], v[jj][5]); o.w = pk_bf16(v[jj][6], v[jj][7]);
                *reinterpret_cast<uint4*>(P.out + (((size_t)n * 2 * P.H + 2 * y + (cb >> 3)) * (2 * P.W) + 2 * x) * 4) = o;
              } else {
                if (use_res) {
#pragma unroll
                  for (int ch = 0; ch < NR; ch++) {
                    const uint4 r = res[jj * NR + ch];
                    const uint32_t rr[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
                    for (int q = 0; q < 4; q++) {
                      const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&rr[q]));
                      const int i = ch * 8 + 2 * q;
                      v[jj][i] = fmaf(v[jj][i], beta_r[i], f.x);
                      v[jj][i + 1] = fmaf(v[jj][i + 1], beta_r[i + 1], f.y);
                    }
                  }
                }
                if (P.act) {
#pragma unroll
                  for (int i = 0; i < CH; i++) v[jj][i] = v[jj][i] > 0.f ? v[jj][i] : v[jj][i] * P.slope;
                }
                bf16* op = P.out + (((size_t)n * P.H + y) * P.W + x) * COUT + co0 + cb;
                if constexpr (CH >= 16) {
#pragma unroll
                  for (int c = 0; c < CH; c += 16) {
                    uint4 o0, o1;
                    o0.x = pk_bf16(v[jj][c], v[jj][c + 1]); o0.y = pk_bf16(v[jj][c + 2], v[jj][c + 3]);
                    o0.z = pk_bf16(v[jj][c + 4], v[jj][c + 5]); o0.w = pk_bf16(v[jj][c + 6], v[jj][c + 7]);
                    o1.x = pk_bf16(v[jj][c + 8], v[jj][c + 9]); o1.y = pk_bf16(v[jj][c + 10], v[jj][c + 11]);
                    o1.z = pk_bf16(v[jj][c + 12], v[jj][c + 13]); o1.w = pk_bf16(v[jj][c + 14], v[jj][c + 15]);
                    st_v8(op + c, o0, o1);
                  }
                } else {
                  uint4 o;
                  o.x = pk_bf16(v[jj][0], v[jj][1]); o.y = pk_bf16(v[jj][2], v[jj][3]);
                  o.z = pk_bf16(v[jj][4], v[jj][5]); o.w = pk_bf16(v[jj][6], v[jj][7]);
                  *reinterpret_cast<uint4*>(op) = o;
                }
              }
            }
          }
        }
      }
    }
  } else {
    // ================= epilogue: 16 warps = 4 TMEM lane groups x NSJ sub-tile classes x NSC channel slices.  A warp
    // keeps its slice's bias / beta in registers, pulls its accumulator columns with one tcgen05.wait::ld, hands the
    // TMEM buffer back to the MMA warp, and only then does the bias / residual / activation math and the stores.
    const int lg = warp & 3;                      // TMEM lane group this warp may access
    const int role = (warp - PW_ - 1) >> 2;     // 0..3
    constexpr int NSJ = KW >= 2 ? 2 : 1;          // sub-tile classes (j mod NSJ)
    constexpr int Q = 4 / NSJ;                    // warps available per lane group and sub-tile class
    // columns per warp: CN / Q when that is 16 or 32, else 16 (32-byte stores; spare warps only hand-shake), 8 only for
    // the Head deconv instance whose 8 columns are one output row
    constexpr int CH = EPI == 1 ? 8 : (CN % Q == 0 && (CN / Q == 16 || CN / Q == 32)) ? CN / Q : 16;
    constexpr int NSC = CN / CH;                  // channel slices in use (<= 4 / NSJ; spare warps only hand-shake)
    static_assert(NSC <= Q && NSC * CH == CN, "epilogue slicing");
    constexpr int NJ = KW / NSJ;                  // sub-tiles per warp and tile
    static_assert(CH == 8 || CH == 16 || CH == 32, "epilogue slice width");
    const int jc = role % NSJ;
    const int cb = (role / NSJ) * CH;
    const bool active = role / NSJ < NSC;
    const int m = lg * 32 + lane;
    const int ly = m >> 3, lx = m & 7;
    const float* bias_c = P.bias_c + co0 + (active ? cb : 0);  // warp-uniform constant-bank reads
    const float* beta_c = P.beta_c + co0 + (active ? cb : 0);
    constexpr bool REGB = TMA && V2_TMA_SLIM;
    float bias_v[REGB ? CH : 1], beta_v[REGB ? CH : 1];
    if constexpr (REGB) {
#pragma unroll
      for (int i = 0; i < CH; i++) { bias_v[i] = bias_c[i]; beta_v[i] = beta_c[i]; }
    }
    const float* bias_r = REGB ? bias_v : bias_c;
    const float* beta_r = REGB ? beta_v : beta_c;
    constexpr int NR = CH / 8;
    const bool tr = warp == PW_ + 1 && lane == 0;
    int it = 0;
    TileWalk tw(tile0, tile_step, P.tiles_x, P.tiles_y);
    StageWalk sw(P.nstages);
    for (; tw.tile < P.num_tiles; tw.next(), sw.next(), it++) {
      const int st = sw.st;
      const int as = it % C::NACC;
      const int tx = tw.tx, ty = tw.ty, n = tw.n;
      const int y = ty * 16 + ly;
      // residual = centre pixels of the input halo tile: pull them into registers as soon as the tile has landed and
      // release the smem stage right away (the stage then only lives for load + MMA, not for the whole epilogue)
      uint4 res[NJ * NR];
      if (tr) V2_TRACE(2, it, 3);  // tile begin
      if (use_res && P.res_global) {
        if (active)
#pragma unroll
        for (int jj = 0; jj < NJ; jj++) {
          const int x = (tx * KW + jj * NSJ + jc) * 8 + lx;
          const uint4* rp = reinterpret_cast<const uint4*>(P.in + (((size_t)n * P.H + y) * P.W + x) * CIN + co0 + cb);
          const bool ok = (y < P.H) && (x < P.W);
#pragma unroll
          for (int ch = 0; ch < NR; ch++) res[jj * NR + ch] = ok ? __ldg(rp + ch) : make_uint4(0u, 0u, 0u, 0u);
        }
      } else if (use_res) {
        if (!(P.ablate & 32)) bar_wait(&full_bar[st], sw.ph);
        if (tr) V2_TRACE(2, it, 4);  // input tile visible
        const uint4* a_st = reinterpret_cast<const uint4*>(a_s + (size_t)st * C::A_BYTES);
        if (active && !(P.ablate & 32))
#pragma unroll
        for (int jj = 0; jj < NJ; jj++) {
          const int pc = (ly + 1) * C::PITCH + 8 * (jj * NSJ + jc) + lx + 1;  // this thread's centre pixel
#pragma unroll
          for (int ch = 0; ch < NR; ch++) {
            if constexpr (PM) {
              const int c8 = (co0 + cb) / 8 + ch;  // 16-byte unit of the pixel's channels
              const uint32_t row = s_u32(a_s + (size_t)st * C::A_BYTES) + (uint32_t)((c8 >> 3) * C::NPIX * 128 + pc * 128);
              const uint32_t ad = row + (uint32_t)(((c8 & 7) ^ ((row >> 7) & 7)) << 4);
              asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];"
                           : "=r"(res[jj * NR + ch].x), "=r"(res[jj * NR + ch].y), "=r"(res[jj * NR + ch].z), "=r"(res[jj * NR + ch].w)
                           : "r"(ad));
            } else {
              res[jj * NR + ch] = a_st[((co0 + cb) / 8 + ch) * C::CS + pc];
            }
          }
        }
        // the stage is refilled by TMA (async proxy): these generic-proxy reads must be ordered before that write.  REQUIRED: without
        // this fence (v2_ablate bit 16, timing experiments only) tools/stress_determinism.py shows results changing from run to run
        // (the mbarrier arrive / wait alone does not order the generic-proxy reads against the TMA write); it costs 4 % of a ResConv launch
        if (TMA && !(P.ablate & 16)) fence_async_smem();
        __syncwarp();
        if (lane == 0) bar_arrive(&empty_bar[st]);
      }
      if (tr) V2_TRACE(2, it, 0);  // residual read, waiting for the accumulator
      if (P.ablate & 1024) {
        if (lane == 0) bar_wait(&tfull_bar[as], (uint32_t)((it / C::NACC) & 1));
        __syncwarp();
      } else if (P.ablate & 2048) {
        uint32_t done;
        do {
          asm volatile("{\n\t.reg .pred P1;\n\tmbarrier.test_wait.parity.shared::cta.b64 P1, [%1], %2;\n\tselp.u32 %0, 1, 0, P1;\n\t}"
                       : "=r"(done) : "r"(s_u32(&tfull_bar[as])), "r"((uint32_t)((it / C::NACC) & 1)) : "memory");
        } while (!done);
      } else
      bar_wait(&tfull_bar[as], (uint32_t)((it / C::NACC) & 1));
      if (tr) V2_TRACE(2, it, 1);  // accumulator complete
      tc_after();
      float v[NJ][CH];
      if (!(P.ablate & 8) && active) {
#pragma unroll
        for (int jj = 0; jj < NJ; jj++)
          ld_cols<CH>(tmem_base + ((uint32_t)(lg * 32) << 16) +
                          (uint32_t)(as * C::ACC_STRIDE + (jj * NSJ + jc) * CN + cb), v[jj]);
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      } else {
#pragma unroll
        for (int jj = 0; jj < NJ; jj++)
#pragma unroll
          for (int i = 0; i < CH; i++) v[jj][i] = 0.f;
      }
      if (tr) V2_TRACE(2, it, 5);  // accumulator in registers
      tc_before();
      __syncwarp();
      if (lane == 0) bar_arrive(&tempty_bar[as]);
      if (tr) V2_TRACE(2, it, 2);  // accumulator released
#pragma unroll
      for (int jj = 0; jj < NJ; jj++) {
        const int j = jj * NSJ + jc;
        const int x = (tx * KW + j) * 8 + lx;
        const bool valid = (y < P.H) && (x < P.W);
        if (valid && active && !(P.ablate & 4)) {
#pragma unroll
          for (int i = 0; i < CH; i++) v[jj][i] += bias_r[i];
          if constexpr (EPI == 1) {
            // Head.cnn3: column = (2p+q)*4 + co; this warp's 8 columns are output row 2y+p (p = cb/8), pixels 2x and
            // 2x+1, 4 channels each
            static_assert(EPI != 1 || (COUT == 16 && CH == 8), "Head deconv instance");
            uint4 o;
            o.x = pk_bf16(v[jj][0], v[jj][1]); o.y = pk_bf16(v[jj][2], v[jj][3]);
            o.z = pk_bf16(v[jj][4], v[jj][5]); o.w = pk_bf16(v[jj][6], v[jj][7]);
            *reinterpret_cast<uint4*>(P.out + (((size_t)n * 2 * P.H + 2 * y + (cb >> 3)) * (2 * P.W) + 2 * x) * 4) = o;
          } else {
            if (use_res && !(P.ablate & 8192)) {
#pragma unroll
              for (int ch = 0; ch < NR; ch++) {
                const uint4 r = res[jj * NR + ch];
                const uint32_t rr[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
                for (int q = 0; q < 4; q++) {
                  const float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&rr[q]));
                  const int i = ch * 8 + 2 * q;
                  v[jj][i] = fmaf(v[jj][i], beta_r[i], f.x);
                  v[jj][i + 1] = fmaf(v[jj][i + 1], beta_r[i + 1], f.y);
                }
              }
            }
            if (P.act && !(P.ablate & 16384)) {
#pragma unroll
              for (int i = 0; i < CH; i++) v[jj][i] = v[jj][i] > 0.f ? v[jj][i] : v[jj][i] * P.slope;
            }
            bf16* op = P.out + (((size_t)n * P.H + y) * P.W + x) * COUT + co0 + cb;
            if ((P.ablate & 4096) && v[jj][0] != 123456.789f) {
              // debug: math kept alive, no stores
            } else if constexpr (CH >= 16) {
#pragma unroll
              for (int c = 0; c < CH; c += 16) {
                uint4 o0, o1;
                o0.x = pk_bf16(v[jj][c], v[jj][c + 1]); o0.y = pk_bf16(v[jj][c + 2], v[jj][c + 3]);
                o0.z = pk_bf16(v[jj][c + 4], v[jj][c + 5]); o0.w = pk_bf16(v[jj][c + 6], v[jj][c + 7]);
                o1.x = pk_bf16(v[jj][c + 8], v[jj][c + 9]); o1.y = pk_bf16(v[jj][c + 10], v[jj][c + 11]);
                o1.z = pk_bf16(v[jj][c + 12], v[jj][c + 13]); o1.w = pk_bf16(v[jj][c + 14], v[jj][c + 15]);
                st_v8(op + c, o0, o1);  // one full 32-byte sector per lane
              }
            } else {
              uint4 o;
              o.x = pk_bf16(v[jj][0], v[jj][1]); o.y = pk_bf16(v[jj][2], v[jj][3]);
              o.z = pk_bf16(v[jj][4], v[jj][5]); o.w = pk_bf16(v[jj][6], v[jj][7]);
              *reinterpret_cast<uint4*>(op) = o;
            }
          }
        }
        if (tr && jj == 0) V2_TRACE(2, it, 6);  // first sub-tile stored
      }
      if (tr) V2_TRACE(2, it, 7);  // tile stored
    }
  }
  tc_before();
  __syncthreads();
  if (warp == PW_) {
    tc_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)C::TMEM_COLS)
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

// instances: Head cnn1/cnn2 (16 ch), block-4 ResConv (32 ch), block-3 ResConv (64 ch)
bool v2_supported(const Layer& L) {
  if (L.kind == LK_DECONV4) return (L.cin == 16 && L.cout == 4) || (L.cin == 32 && L.cout == 52);  // Head.cnn3, final lastconv
  if (L.kind != LK_CONV3 || L.stride != 1 || L.cin != L.cout || L.cin != L.cin_pad) return false;
  return L.cin == 16 || L.cin == 32 || L.cin == 64 || L.cin == 96 || L.cin == 128;
}
static int v2_nsplit(int c) { return c == 96 ? 2 : c == 128 ? 4 : 1; }

// Packs Conv2d weights [Co][Ci][3][3] into the resident smem image [tap][Ci/8][Co][8] bf16.
int v2_build_layer(vfi_handle* h, Layer& L, const std::vector<float>& w) {
  if (!v2_supported(L)) return 0;
  const int nch = L.cin / 8;
  if (L.kind == LK_DECONV4 && L.cout == 52) {
    // final block's lastconv, channels 0-19 only: column (2p+q)*20 + (2i+jj)*5 + c <- deconv channel 4c + 2i + jj of
    // phase (p,q); kernel element ky = p + 1 - 2*dy, kx = q + 1 - 2*dx of tap (dy,dx); w: [Ci][52][4][4]
    std::vector<uint16_t> pk((size_t)9 * 32 * 80, 0);
    std::vector<float> bias_h(L.cout);
    VFI_CUDA(h, cudaMemcpy(bias_h.data(), L.bias, L.cout * sizeof(float), cudaMemcpyDeviceToHost));
    for (int i = 0; i < 128; i++) { L.h_bias_v2[i] = 0.f; L.h_beta[i] = 1.f; }
    for (int ph = 0; ph < 4; ph++)
      for (int k = 0; k < 4; k++)
        for (int c = 0; c < 5; c++) {
          const int col = ph * 20 + k * 5 + c, dc = 4 * c + k;
          L.h_bias_v2[col] = bias_h[dc];
          for (int tap = 0; tap < 9; tap++) {
            const int dy = tap / 3 - 1, dx = tap % 3 - 1;
            const int ky = (ph >> 1) + 1 - 2 * dy, kx = (ph & 1) + 1 - 2 * dx;
            if (ky < 0 || ky > 3 || kx < 0 || kx > 3) continue;
            for (int ci = 0; ci < 32; ci++)
              pk[(((size_t)tap * nch + ci / 8) * 80 + col) * 8 + ci % 8] =
                  f2bf_v2(w[(((size_t)ci * 52 + dc) * 4 + ky) * 4 + kx]);
          }
        }
    VFI_CUDA(h, cudaMalloc(&L.wv2, pk.size() * 2));
    VFI_CUDA(h, cudaMemcpy(L.wv2, pk.data(), pk.size() * 2, cudaMemcpyHostToDevice));
    return 0;
  }
  if (L.kind == LK_DECONV4) {
    // ConvTranspose2d(16,4,4,2,1): output pixel (2y+p, 2x+q) reads input (y+dy, x+dx) through kernel element
    // ky = p + 1 - 2*dy, kx = q + 1 - 2*dx.  Column = (2p+q)*4 + co of a 3x3 "conv" over the input grid; taps a
    // phase does not use are zero.  w: [Ci][Co][4][4].
    std::vector<uint16_t> pk((size_t)9 * 16 * 16, 0);
    for (int tap = 0; tap < 9; tap++) {
      const int dy = tap / 3 - 1, dx = tap % 3 - 1;
      for (int ph = 0; ph < 4; ph++) {
        const int ky = (ph >> 1) + 1 - 2 * dy, kx = (ph & 1) + 1 - 2 * dx;
        if (ky < 0 || ky > 3 || kx < 0 || kx > 3) continue;
        for (int co = 0; co < 4; co++)
          for (int ci = 0; ci < 16; ci++)
            pk[(((size_t)tap * nch + ci / 8) * 16 + ph * 4 + co) * 8 + ci % 8] =
                f2bf_v2(w[(((size_t)ci * 4 + co) * 4 + ky) * 4 + kx]);
      }
    }
    VFI_CUDA(h, cudaMalloc(&L.wv2, pk.size() * 2));
    VFI_CUDA(h, cudaMemcpy(L.wv2, pk.data(), pk.size() * 2, cudaMemcpyHostToDevice));
    float b16[16], bh[4];
    VFI_CUDA(h, cudaMemcpy(bh, L.bias, sizeof(bh), cudaMemcpyDeviceToHost));
    for (int i = 0; i < 16; i++) b16[i] = bh[i & 3];
    for (int i = 0; i < 128; i++) { L.h_bias_v2[i] = b16[i & 15]; L.h_beta[i] = 1.f; }
    VFI_CUDA(h, cudaMalloc(&L.bias_v2, sizeof(b16)));
    VFI_CUDA(h, cudaMemcpy(L.bias_v2, b16, sizeof(b16), cudaMemcpyHostToDevice));
    return 0;
  }
  VFI_CUDA(h, cudaMemcpy(L.h_bias_v2, L.bias, L.cout * sizeof(float), cudaMemcpyDeviceToHost));
  if (L.has_beta) VFI_CUDA(h, cudaMemcpy(L.h_beta, L.beta, L.cout * sizeof(float), cudaMemcpyDeviceToHost));
  // [split][tap][Ci/8][Co/nsplit][8]: each output-channel group's slice is one contiguous resident image
  const int nsp = v2_nsplit(L.cin), cn = L.cout / nsp;
  std::vector<uint16_t> pk((size_t)9 * L.cin * L.cout);
  for (int tap = 0; tap < 9; tap++)
    for (int j = 0; j < nch; j++)
      for (int co = 0; co < L.cout; co++)
        for (int e = 0; e < 8; e++) {
          const int ci = j * 8 + e;
          pk[((((size_t)(co / cn) * 9 + tap) * nch + j) * cn + co % cn) * 8 + e] =
              f2bf_v2(w[((size_t)co * L.cin + ci) * 9 + tap]);
        }
  VFI_CUDA(h, cudaMalloc(&L.wv2, pk.size() * 2));
  VFI_CUDA(h, cudaMemcpy(L.wv2, pk.data(), pk.size() * 2, cudaMemcpyHostToDevice));
  return 0;
}

template <int CIN, int COUT, int KW, int EPI, int NSPLIT = 1, bool TMA = false, bool PM = false>
static int launch_v2(vfi_handle* h, V2Params& P, cudaStream_t s) {
  using C = V2Cfg<CIN, COUT / NSPLIT, KW, TMA, PM>;
  CUtensorMap tmA;
  memset(&tmA, 0, sizeof(tmA));
  if (TMA) {
    if (PM) { if (int rc = tc_make_tmap_sw128(h, &tmA, P.in, CIN, P.W, P.H, P.N, C::PITCH, 18)) return rc; }
    else if (int rc = tc_make_tmap_plain(h, &tmA, P.in, CIN, P.W, P.H, P.N, 8, C::PITCH, 18)) return rc;
  }
  const size_t fixed = (size_t)C::W_BYTES + C::BB_BYTES;
  const size_t budget = 220 * 1024 - fixed;
  P.nstages = (int)std::min<size_t>(V2_MAX_STAGES, budget / C::A_BYTES);
  if (P.nstages < 2) VFI_FAIL(h, VFI_E_ARG, "conv_v2: shared memory budget (Cin=%d)", CIN);
  P.tiles_x = cdiv(P.W, 8 * KW);
  P.tiles_y = cdiv(P.H, 16);
  P.num_tiles = P.N * P.tiles_x * P.tiles_y;
  const size_t smem = fixed + (size_t)P.nstages * C::A_BYTES;
  VFI_CUDA(h, cudaFuncSetAttribute(conv_v2_kernel<CIN, COUT, KW, EPI, NSPLIT, TMA, PM>,
                                   cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024));
  const int grid = std::min(P.num_tiles, h->sm_count / NSPLIT) * NSPLIT;
  conv_v2_kernel<CIN, COUT, KW, EPI, NSPLIT, TMA, PM><<<grid, (TMA && V2_TMA_SLIM) ? 32 + 32 + 32 * V2_EPI_WARPS : V2_THREADS, smem, s>>>(P, tmA);
  VFI_LAUNCHED(h);
  return 0;
}

int k_conv_v2(vfi_handle* h, const Layer& L, const bf16* in, bf16* out, int N, int H, int W, float slope,
              cudaStream_t s) {
  ProfScope _ps(h, PC_CONV_TC, s);
  if (!L.wv2) VFI_FAIL(h, VFI_E_STATE, "conv_v2: layer not packed");
  V2Params P;
  memset(&P, 0, sizeof(P));
  P.in = in; P.out = out; P.wpack = L.wv2; P.bias = L.bias; P.beta = L.has_beta ? L.beta : nullptr;
  P.N = N; P.H = H; P.W = W;
  P.act = L.act ? 1 : 0;
  P.slope = slope;
  P.ablate = h->opt_v2_ablate;
  P.res_global = h->opt_v2_res_global;
  for (int i = 0; i < 128; i++) { P.bias_c[i] = L.h_bias_v2[i]; P.beta_c[i] = L.has_beta ? L.h_beta[i] : 1.f; }
  if (h->opt_v2_trace) {
    // debug: time-stamp CTA 0's producer / MMA / epilogue hand-shakes of this launch and dump them for tools/v2_trace.py
    static unsigned long long* tbuf = nullptr;
    const size_t tbytes = 3 * 64 * 8 * sizeof(unsigned long long);
    if (!tbuf) VFI_CUDA(h, cudaMalloc(&tbuf, tbytes));
    VFI_CUDA(h, cudaMemsetAsync(tbuf, 0, tbytes, s));
    P.trace = tbuf;
    int rc;
    switch (L.cin) {
      case 16: rc = launch_v2<16, 16, 4, 0>(h, P, s); break;
      case 32: rc = h->opt_v2_tma ? launch_v2<32, 32, 4, 0, 1, true>(h, P, s) : launch_v2<32, 32, 4, 0>(h, P, s); break;
      default: rc = h->opt_v2_tma ? launch_v2<64, 64, 2, 0, 1, true>(h, P, s) : launch_v2<64, 64, 2, 0>(h, P, s); break;
    }
    if (rc) return rc;
    std::vector<unsigned long long> host(3 * 64 * 8);
    VFI_CUDA(h, cudaStreamSynchronize(s));
    VFI_CUDA(h, cudaMemcpy(host.data(), tbuf, tbytes, cudaMemcpyDeviceToHost));
    char path[128];
    snprintf(path, sizeof(path), "gpurun_out/v2_trace_c%d.bin", L.cin);
    if (FILE* f = fopen(path, "wb")) { fwrite(host.data(), 1, tbytes, f); fclose(f); }
    return 0;
  }
  if (L.kind == LK_DECONV4 && L.cout == 52) VFI_FAIL(h, VFI_E_ARG, "conv_v2: use k_conv_v2_last for the final lastconv");
  if (L.kind == LK_DECONV4) {
    P.bias = L.bias_v2; P.beta = nullptr; P.act = 0;
    return h->opt_v2_tma ? launch_v2<16, 16, 4, 1, 1, true>(h, P, s) : launch_v2<16, 16, 4, 1>(h, P, s);
  }
  switch (L.cin) {
    case 16: return h->opt_v2_tma ? launch_v2<16, 16, 4, 0, 1, true>(h, P, s) : launch_v2<16, 16, 4, 0>(h, P, s);
    case 32: return h->opt_v2_tma ? launch_v2<32, 32, 4, 0, 1, true>(h, P, s) : launch_v2<32, 32, 4, 0>(h, P, s);
    // KW = 2: 3 stages beside the 72 KB of weights
    case 64: return (h->opt_v2_tma && h->opt_v2_pm) ? launch_v2<64, 64, 2, 0, 1, true, true>(h, P, s)
           : h->opt_v2_tma ? launch_v2<64, 64, 2, 0, 1, true>(h, P, s) : launch_v2<64, 64, 2, 0>(h, P, s);
    // 2 x 48 output channels: 81 KB of weights per CTA
    case 96: return h->opt_v2_tma ? launch_v2<96, 96, 1, 0, 2, true>(h, P, s) : launch_v2<96, 96, 1, 0, 2>(h, P, s);
    // 4 x 32 output channels: 72 KB of weights per CTA
    case 128: return h->opt_v2_tma ? launch_v2<128, 128, 1, 0, 4, true>(h, P, s) : launch_v2<128, 128, 1, 0, 4>(h, P, s);
  }
  VFI_FAIL(h, VFI_E_ARG, "conv_v2: no instance for Cin=%d", L.cin);
}

// Final block's lastconv (ConvTranspose 4x4 s2 + PixelShuffle(2), output channels 0-4 = flow delta + mask):
// in [N, H, W, 32] bf16 -> low [N, 4H, 4W, low_c] fp32 (low_c = 8 in the network, 16 in the vfi_debug_conv layout)
int k_conv_v2_last(vfi_handle* h, const Layer& L, const bf16* in, float* low, int low_c, int N, int H, int W,
                   cudaStream_t s) {
  ProfScope _ps(h, PC_CONV_TC, s);
  if (!L.wv2 || L.kind != LK_DECONV4 || L.cout != 52) VFI_FAIL(h, VFI_E_STATE, "conv_v2_last: layer not packed");
  V2Params P;
  memset(&P, 0, sizeof(P));
  P.in = in; P.wpack = L.wv2; P.bias = L.bias; P.beta = nullptr;
  P.low_out = low; P.low_c = low_c;
  P.N = N; P.H = H; P.W = W;
  for (int i = 0; i < 128; i++) { P.bias_c[i] = L.h_bias_v2[i]; P.beta_c[i] = 1.f; }
  return h->opt_v2_tma ? launch_v2<32, 80, 2, 2, 1, true>(h, P, s) : launch_v2<32, 80, 2, 2>(h, P, s);
}
