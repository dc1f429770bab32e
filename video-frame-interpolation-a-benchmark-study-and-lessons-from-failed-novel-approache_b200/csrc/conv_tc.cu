// conv_tc.cu -- IFNet convolutions as implicit GEMM on the 5th-gen tensor cores (sm_100a).
//
//   D[128 pixels x N] (fp32, TMEM) += A[128 pixels x Cblk] (bf16, smem via TMA) * B[N x Cblk]^T (bf16, smem via TMA)
//
// One k-step = one filter tap x one channel chunk.  The A tile of a tap is a 4-D TMA box
// [Cblk, tw, th, 1] of the NHWC activation tensor shifted by the tap offset; out-of-bounds rows
// and columns are zero-filled by TMA, which IS the convolution's zero padding.  Stride-2 convs read
// the four phase planes their producer wrote (plane = parity of the input row/col), so every tap is
// still a dense stride-1 box.  ConvTranspose2d(4,2,1) is 4 output phases x (2x2 taps): the 9 distinct
// input taps are each loaded ONCE and feed the accumulators of every phase that uses them (up to
// 256 TMEM columns), followed by a PixelShuffle(2) epilogue.
//
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM allocator + single-thread
// tcgen05.mma issuer, warps 2..5 = epilogue (tcgen05.ld -> bias / beta*x+res / LeakyReLU -> global).
// Persistent CTAs, multi-stage smem ring (full/empty mbarriers), double-buffered TMEM accumulator.
// Layer semantics: SURVEY.md Appendix A.2 (conv(), ResConv, IFBlock.lastconv).
#include <algorithm>
#include <map>

#include "common.cuh"
#include "umma.cuh"

// ---------------------------------------------------------------------------------------------
struct TcParams {
  int N, Ho, Wo;  // M-space pixel grid
  int th, tw, tiles_x, tiles_y, num_tiles;
  int planes;
  int nsteps, cblk, ksub;
  int a_bytes, stage_bytes, nstages;
  int n_total, acc_stride, b_box, tmem_cols;
  uint32_t sbo16, layout_type;
  int epi_mode;  // 0 conv->plain bf16, 1 conv->phase-plane bf16, 2 lastconv->pixel-shuffled fp32 low13
  int cout, act;
  bf16* low_mf;  // epi_mode 2, non-null: LowPlanes output (out = flow plane)
  bf16* low_f7;
  int low_c;     // epi_mode 2: floats per output texel, 16 (13 channels) or 8 (5 channels: flow delta + mask)
  float slope;
  const float* bias;
  const float* beta;
  const bf16* res;
  void* out;
  TcStep steps[VFI_MAX_STEPS];
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, uint32_t cnt) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(cnt) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred P1;\n\t"
      "WAIT_LOOP:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t"
      "@P1 bra WAIT_DONE;\n\t"
      "bra WAIT_LOOP;\n\t"
      "WAIT_DONE:\n\t}" ::"r"(smem_u32(b)), "r"(parity)
      : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* b) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(b)) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* b, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_load_4d(const CUtensorMap* m, uint64_t* bar, void* dst, int c0, int c1, int c2,
                                            int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(smem_u32(dst)), "l"((uint64_t)m), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void tma_load_2d(const CUtensorMap* m, uint64_t* bar, void* dst, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(dst)), "l"((uint64_t)m), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void tc_mma_bf16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                            uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float* v) {
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
// K-major swizzled smem matrix descriptor (cute::UMMA::SmemDescriptor bit layout, sm_100 version = 1)
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t sbo16, uint32_t layout_type) {
  uint64_t d = (uint64_t)((saddr & 0x3FFFF) >> 4);
  d |= (uint64_t)1 << 16;               // LBO (ignored for swizzled K-major)
  d |= (uint64_t)(sbo16 & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;               // descriptor version
  d |= (uint64_t)(layout_type & 7) << 61;
  return d;
}
__device__ __forceinline__ uint32_t make_idesc(int n) {
  // c_format F32 (1<<4), a/b format BF16 (1<<7, 1<<10), K-major A and B, N>>3 at bit 17, M=128 (8) at bit 24
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | (8u << 24);
}
__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  __nv_bfloat162 t = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&t);
}

#define TC_THREADS 192
#define TC_MAX_STAGES 8

__global__ void __launch_bounds__(TC_THREADS, 1)
conv_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const __grid_constant__ TcParams P) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  __shared__ uint64_t full_bar[TC_MAX_STAGES], empty_bar[TC_MAX_STAGES], tfull_bar[2], tempty_bar[2];
  __shared__ uint32_t tmem_base_smem;

  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int i = 0; i < P.nstages; i++) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
    for (int i = 0; i < 2; i++) { mbar_init(&tfull_bar[i], 1); mbar_init(&tempty_bar[i], 4); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"((uint64_t)&tmA) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"((uint64_t)&tmB) : "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_smem)),
                 "r"((uint32_t)P.tmem_cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;

  if (warp == 0) {
    // ================= TMA producer =================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      umma::TileWalk tw((int)blockIdx.x, (int)gridDim.x, P.tiles_x, P.tiles_y);   // no per-tile integer divisions (umma.cuh)
      for (; tw.tile < P.num_tiles; tw.next()) {
        const int tx = tw.tx, ty = tw.ty, n = tw.n;
        int x0 = tx * P.tw, y0 = ty * P.th;
        for (int s = 0; s < P.nsteps; s++) {
          const TcStep& st = P.steps[s];
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + (size_t)stage * P.stage_bytes;
          uint8_t* sb = sa + P.a_bytes;
          mbar_expect_tx(&full_bar[stage], (uint32_t)(P.a_bytes + st.b_rows * P.cblk * 2));
          tma_load_4d(&tmA, &full_bar[stage], sa, st.c0, x0 + st.dx, y0 + st.dy, n * P.planes + st.plane);
          for (int r = 0; r < st.b_rows; r += P.b_box)
            tma_load_2d(&tmB, &full_bar[stage], sb + (size_t)r * P.cblk * 2, 0, st.b_row + r);
          if (++stage == P.nstages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer (one thread) =================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int tile = blockIdx.x; tile < P.num_tiles; tile += gridDim.x, it++) {
        const int as = it & 1;
        const uint32_t aphase = (it >> 1) & 1;
        mbar_wait(&tempty_bar[as], aphase ^ 1);
        tc_fence_after();
        const uint32_t dcol = tmem_base + (uint32_t)(as * P.acc_stride);
        for (int s = 0; s < P.nsteps; s++) {
          const TcStep& st = P.steps[s];
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + (size_t)stage * P.stage_bytes);
          const uint32_t sb = sa + P.a_bytes;
          for (int g = 0; g < st.nseg; g++) {
            const TcSeg sg = st.seg[g];
            const uint32_t idesc = make_idesc(sg.n);
            const uint64_t ad = make_desc(sa, P.sbo16, P.layout_type);
            const uint64_t bd = make_desc(sb + (uint32_t)sg.b_row_off * P.cblk * 2, P.sbo16, P.layout_type);
            for (int k = 0; k < P.ksub; k++)
              tc_mma_bf16(dcol + sg.tmem_col, ad + (uint64_t)(2 * k), bd + (uint64_t)(2 * k), idesc,
                          (sg.accumulate || k > 0) ? 1u : 0u);
          }
          tc_commit(&empty_bar[stage]);  // frees the smem slot once these MMAs have read it
          if (++stage == P.nstages) { stage = 0; phase ^= 1; }
        }
        tc_commit(&tfull_bar[as]);  // accumulator complete -> epilogue
      }
    }
  } else {
    // ================= epilogue warps =================
    const int lg = warp & 3;  // TMEM lane group this warp may access
    const int m = lg * 32 + lane;
    int it = 0;
    const int my = m / P.tw, mx = m % P.tw;   // this thread's pixel inside a tile: loop invariant
    umma::TileWalk tw((int)blockIdx.x, (int)gridDim.x, P.tiles_x, P.tiles_y);
    for (; tw.tile < P.num_tiles; tw.next(), it++) {
      const int as = it & 1;
      const uint32_t aphase = (it >> 1) & 1;
      const int tx = tw.tx, ty = tw.ty, n = tw.n;
      const int y = ty * P.th + my, x = tx * P.tw + mx;
      const bool valid = (y < P.Ho) && (x < P.Wo);
      mbar_wait(&tfull_bar[as], aphase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(lg * 32) << 16) + (uint32_t)(as * P.acc_stride);
      if (P.epi_mode <= 1) {
        size_t opix;
        if (P.epi_mode == 0) opix = ((size_t)n * P.Ho + y) * P.Wo + x;
        else opix = ((((size_t)n * 4 + (y & 1) * 2 + (x & 1)) * (P.Ho >> 1) + (y >> 1)) * (P.Wo >> 1) + (x >> 1));
        bf16* op = reinterpret_cast<bf16*>(P.out) + opix * P.cout;
        const bf16* rp = P.res ? P.res + (((size_t)n * P.Ho + y) * P.Wo + x) * P.cout : nullptr;
        for (int c = 0; c < P.cout; c += 16) {
          float v[16];
          tmem_ld16(taddr + c, v);
          if (valid) {
#pragma unroll
            for (int i = 0; i < 16; i++) v[i] += __ldg(P.bias + c + i);
            if (P.beta) {
              uint4 r0 = __ldg(reinterpret_cast<const uint4*>(rp + c));
              uint4 r1 = __ldg(reinterpret_cast<const uint4*>(rp + c + 8));
              const uint32_t rr[8] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
#pragma unroll
              for (int i = 0; i < 8; i++) {
                float2 f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&rr[i]));
                v[2 * i] = v[2 * i] * __ldg(P.beta + c + 2 * i) + f.x;
                v[2 * i + 1] = v[2 * i + 1] * __ldg(P.beta + c + 2 * i + 1) + f.y;
              }
            }
            if (P.act) {
#pragma unroll
              for (int i = 0; i < 16; i++) v[i] = v[i] > 0.f ? v[i] : v[i] * P.slope;
            }
            uint4 o0, o1;
            o0.x = pack_bf16(v[0], v[1]); o0.y = pack_bf16(v[2], v[3]); o0.z = pack_bf16(v[4], v[5]); o0.w = pack_bf16(v[6], v[7]);
            o1.x = pack_bf16(v[8], v[9]); o1.y = pack_bf16(v[10], v[11]); o1.z = pack_bf16(v[12], v[13]); o1.w = pack_bf16(v[14], v[15]);
            st_v8(op + c, o0, o1);
          }
        }
      } else {
        // lastconv: 4 phases x 64 columns (52 used) -> PixelShuffle(2) -> low [N, 4Ho, 4Wo, low_c] fp32
        float* ob = reinterpret_cast<float*>(P.out);
        const int H4 = 4 * P.Ho, W4 = 4 * P.Wo;
        for (int ph = 0; ph < 4; ph++) {
          const int p = ph >> 1, q = ph & 1;
          if (P.low_mf) {
            // bf16 planes: the two texels (jj = 0, 1) of a row pair up into full 32-byte sectors of the flow and the
            // mask/feat plane
            float v[64];
#pragma unroll
            for (int c = 0; c < 64; c += 16) tmem_ld16(taddr + ph * 64 + c, v + c);
            if (valid) {
#pragma unroll
              for (int i = 0; i < 2; i++) {
                float o[2][13];
#pragma unroll
                for (int jj = 0; jj < 2; jj++)
#pragma unroll
                  for (int cc = 0; cc < 13; cc++) o[jj][cc] = v[cc * 4 + 2 * i + jj] + __ldg(P.bias + cc * 4 + 2 * i + jj);
                const size_t t0 = ((size_t)n * H4 + 4 * y + 2 * p + i) * W4 + 4 * x + 2 * q;
                st_v8f(ob + t0 * 4, make_float4(o[0][0], o[0][1], o[0][2], o[0][3]), make_float4(o[1][0], o[1][1], o[1][2], o[1][3]));
                uint4 m0, m1;
                m0.x = pack_bf16(o[0][4], o[0][5]); m0.y = pack_bf16(o[0][6], o[0][7]); m0.z = pack_bf16(o[0][8], o[0][9]); m0.w = pack_bf16(o[0][10], o[0][11]);
                m1.x = pack_bf16(o[1][4], o[1][5]); m1.y = pack_bf16(o[1][6], o[1][7]); m1.z = pack_bf16(o[1][8], o[1][9]); m1.w = pack_bf16(o[1][10], o[1][11]);
                st_v8(P.low_mf + t0 * 8, m0, m1);
                *reinterpret_cast<uint32_t*>(P.low_f7 + t0) = pack_bf16(o[0][12], o[1][12]);
              }
            }
          } else if (P.low_c == 16) {
            float v[64];
#pragma unroll
            for (int c = 0; c < 64; c += 16) tmem_ld16(taddr + ph * 64 + c, v + c);
            if (valid) {
#pragma unroll
              for (int k = 0; k < 4; k++) {
                const int i = k >> 1, jj = k & 1;
                float o[16];
#pragma unroll
                for (int cc = 0; cc < 13; cc++) o[cc] = v[cc * 4 + k] + __ldg(P.bias + cc * 4 + k);
                o[13] = o[14] = o[15] = 0.f;
                float4* dst = reinterpret_cast<float4*>(
                    ob + (((size_t)n * H4 + 4 * y + 2 * p + i) * W4 + 4 * x + 2 * q + jj) * 16);
                st_v8f(reinterpret_cast<float*>(dst), make_float4(o[0], o[1], o[2], o[3]), make_float4(o[4], o[5], o[6], o[7]));
                st_v8f(reinterpret_cast<float*>(dst + 2), make_float4(o[8], o[9], o[10], o[11]),
                       make_float4(o[12], o[13], o[14], o[15]));
              }
            }
          } else {
            // final block: channels 0..4 (deconv channels 0..19) only
            float v[32];
            tmem_ld16(taddr + ph * 64, v);
            tmem_ld16(taddr + ph * 64 + 16, v + 16);
            if (valid) {
#pragma unroll
              for (int k = 0; k < 4; k++) {
                const int i = k >> 1, jj = k & 1;
                float o[5];
#pragma unroll
                for (int cc = 0; cc < 5; cc++) o[cc] = v[cc * 4 + k] + __ldg(P.bias + cc * 4 + k);
                float4* dst = reinterpret_cast<float4*>(
                    ob + (((size_t)n * H4 + 4 * y + 2 * p + i) * W4 + 4 * x + 2 * q + jj) * 8);
                st_v8f(reinterpret_cast<float*>(dst), make_float4(o[0], o[1], o[2], o[3]), make_float4(o[4], 0.f, 0.f, 0.f));
              }
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty_bar[as]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)P.tmem_cols)
                 : "memory");
  }
}

// =============================================================================================
// host side
// =============================================================================================
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static int get_encode(vfi_handle* h, PFN_encodeTiled* fn) {
  if (!h->tmap_encode) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qr;
    VFI_CUDA(h, cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qr));
    if (!p || qr != cudaDriverEntryPointSuccess) VFI_FAIL(h, VFI_E_CUDA, "cuTensorMapEncodeTiled not available");
    h->tmap_encode = p;
  }
  *fn = (PFN_encodeTiled)h->tmap_encode;
  return 0;
}

static CUtensorMapSwizzle swizzle_for(int cblk) {
  return cblk == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : cblk == 32 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B;
}

static uint16_t f2bf(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffff) > 0x7f800000) return 0x7fc0;
  u += 0x7fff + ((u >> 16) & 1);  // round to nearest even
  return (uint16_t)(u >> 16);
}

// Packs the layer's weights into the K-major [rows][cblk] bf16 matrix the B-operand TMA reads and
// builds the k-step table.  w_torch: Conv2d [Co][Ci][3][3] or ConvTranspose2d [Ci][Co][4][4].
int tc_build_layer(vfi_handle* h, Layer& L, const std::vector<float>& w) {
  const int cp = L.cin_pad;
  L.cblk = (cp % 64 == 0) ? 64 : (cp % 32 == 0) ? 32 : 16;
  const int nch = cp / L.cblk;
  std::vector<uint16_t> pk;
  std::vector<float> bias_tc;
  int ns = 0;
  if (L.kind == LK_CONV3) {
    L.n_total = L.cout;
    L.b_box = L.cout;
    if (9 * nch > VFI_MAX_STEPS || (L.cout % 16) || L.cout > 256) VFI_FAIL(h, VFI_E_ARG, "tc layer shape unsupported");
    pk.assign((size_t)9 * nch * L.cout * L.cblk, 0);
    for (int r = 0; r < 3; r++)
      for (int q = 0; q < 3; q++)
        for (int ch = 0; ch < nch; ch++) {
          TcStep& st = L.steps[ns];
          memset(&st, 0, sizeof(st));
          if (L.stride == 1) { st.dy = r - 1; st.dx = q - 1; st.plane = 0; }
          else {
            int pr = (r == 1) ? 0 : 1, pq = (q == 1) ? 0 : 1;
            st.dy = (r == 0) ? -1 : 0; st.dx = (q == 0) ? -1 : 0; st.plane = pr * 2 + pq;
          }
          st.c0 = ch * L.cblk;
          st.b_row = ns * L.cout;
          st.b_rows = L.cout;
          st.nseg = 1;
          st.seg[0].b_row_off = 0; st.seg[0].n = L.cout; st.seg[0].tmem_col = 0; st.seg[0].accumulate = ns > 0;
          for (int co = 0; co < L.cout; co++)
            for (int k = 0; k < L.cblk; k++) {
              int ci = st.c0 + k;
              float v = ci < L.cin ? w[(((size_t)co * L.cin + ci) * 3 + r) * 3 + q] : 0.f;
              pk[((size_t)st.b_row + co) * L.cblk + k] = f2bf(v);
            }
          ns++;
        }
    bias_tc.assign(L.cout, 0.f);
  } else {
    // ConvTranspose2d(c, 52, 4, 2, 1): phase (p,q) of output pixel (2y+p, 2x+q) reads input (y+dy, x+dx) with
    // ky = p + 1 - 2*dy, kx = q + 1 - 2*dx.  Phase ph = 2p+q owns accumulator columns [64*ph, 64*ph+52).
    if (L.cout > 64 || 9 * nch > VFI_MAX_STEPS) VFI_FAIL(h, VFI_E_ARG, "tc deconv shape unsupported");
    L.n_total = 256;
    L.b_box = 64;
    const int order[9][2] = {{0, 0}, {-1, -1}, {-1, 0}, {-1, 1}, {0, -1}, {0, 1}, {1, -1}, {1, 0}, {1, 1}};
    int row = 0;
    for (int t = 0; t < 9; t++) {
      int dy = order[t][0], dx = order[t][1];
      int plist[2], qlist[2], np_ = 0, nq = 0;
      if (dy <= 0) plist[np_++] = 0;
      if (dy >= 0) plist[np_++] = 1;
      if (dx <= 0) qlist[nq++] = 0;
      if (dx >= 0) qlist[nq++] = 1;
      int phases[4], nph = 0;
      for (int a = 0; a < np_; a++)
        for (int b = 0; b < nq; b++) phases[nph++] = plist[a] * 2 + qlist[b];
      std::sort(phases, phases + nph);
      for (int ch = 0; ch < nch; ch++) {
        TcStep& st = L.steps[ns];
        memset(&st, 0, sizeof(st));
        st.dy = dy; st.dx = dx; st.plane = 0; st.c0 = ch * L.cblk;
        st.b_row = row; st.b_rows = 64 * nph;
        // contiguous runs of phases -> MMA segments
        int g = 0, i = 0;
        while (i < nph) {
          int j = i;
          while (j + 1 < nph && phases[j + 1] == phases[j] + 1) j++;
          st.seg[g].b_row_off = 64 * i; st.seg[g].n = 64 * (j - i + 1); st.seg[g].tmem_col = 64 * phases[i];
          st.seg[g].accumulate = ns > 0;
          g++; i = j + 1;
        }
        st.nseg = g;
        pk.resize((size_t)(row + 64 * nph) * L.cblk, 0);
        for (int a = 0; a < nph; a++) {
          int p = phases[a] >> 1, q = phases[a] & 1;
          int ky = p + 1 - 2 * dy, kx = q + 1 - 2 * dx;
          for (int co = 0; co < L.cout; co++)
            for (int k = 0; k < L.cblk; k++) {
              int ci = st.c0 + k;
              float v = ci < L.cin ? w[(((size_t)ci * L.cout + co) * 4 + ky) * 4 + kx] : 0.f;
              pk[((size_t)row + 64 * a + co) * L.cblk + k] = f2bf(v);
            }
        }
        row += 64 * nph;
        ns++;
      }
    }
  }
  L.nsteps = ns;
  L.rows_total = (int)(pk.size() / L.cblk);
  VFI_CUDA(h, cudaMalloc(&L.wtc, pk.size() * 2));
  VFI_CUDA(h, cudaMemcpy(L.wtc, pk.data(), pk.size() * 2, cudaMemcpyHostToDevice));
  return 0;
}

struct TmapKey {
  const void* p; int d0, d1, d2, d3, b0, b1, b2, rank;
  bool operator<(const TmapKey& o) const { return memcmp(this, &o, sizeof(TmapKey)) < 0; }
};
static std::map<TmapKey, CUtensorMap>& tmap_cache() { static std::map<TmapKey, CUtensorMap> c; return c; }

static int make_tmap_act(vfi_handle* h, CUtensorMap* out, const bf16* base, int C, int W, int H, int NP, int cblk,
                         int tw, int th) {
  TmapKey key;
  memset(&key, 0, sizeof(key));
  key.p = base; key.d0 = C; key.d1 = W; key.d2 = H; key.d3 = NP; key.b0 = cblk; key.b1 = tw; key.b2 = th; key.rank = 4;
  auto it = tmap_cache().find(key);
  if (it != tmap_cache().end()) { *out = it->second; return 0; }
  PFN_encodeTiled enc;
  if (int rc = get_encode(h, &enc)) return rc;
  cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)NP};
  cuuint64_t strides[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
  cuuint32_t box[4] = {(cuuint32_t)cblk, (cuuint32_t)tw, (cuuint32_t)th, 1};
  cuuint32_t es[4] = {1, 1, 1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, (void*)base, dims, strides, box, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle_for(cblk), CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) VFI_FAIL(h, VFI_E_CUDA, "cuTensorMapEncodeTiled(act C=%d W=%d H=%d NP=%d) failed: %d", C, W, H, NP, (int)r);
  tmap_cache()[key] = *out;
  return 0;
}
int tc_make_tmap_sw128(vfi_handle* h, void* out_tmap, const bf16* base, int C, int W, int H, int N, int bw, int bh) {
  return make_tmap_act(h, reinterpret_cast<CUtensorMap*>(out_tmap), base, C, W, H, N, 64, bw, bh);
}
int tc_make_tmap_plain(vfi_handle* h, void* out_tmap, const bf16* base, int C, int W, int H, int N, int bc, int bw, int bh) {
  CUtensorMap* out = reinterpret_cast<CUtensorMap*>(out_tmap);
  TmapKey key;
  memset(&key, 0, sizeof(key));
  key.p = base; key.d0 = C; key.d1 = W; key.d2 = H; key.d3 = N; key.b0 = bc; key.b1 = bw; key.b2 = bh; key.rank = 14;  // 14: plain
  auto it = tmap_cache().find(key);
  if (it != tmap_cache().end()) { *out = it->second; return 0; }
  PFN_encodeTiled enc;
  if (int rc = get_encode(h, &enc)) return rc;
  cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
  cuuint64_t strides[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
  cuuint32_t box[4] = {(cuuint32_t)bc, (cuuint32_t)bw, (cuuint32_t)bh, 1};
  cuuint32_t es[4] = {1, 1, 1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, (void*)base, dims, strides, box, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) VFI_FAIL(h, VFI_E_CUDA, "cuTensorMapEncodeTiled(plain C=%d W=%d H=%d N=%d box %dx%dx%d) failed: %d", C, W, H, N, bc, bw, bh, (int)r);
  tmap_cache()[key] = *out;
  return 0;
}
static int make_tmap_w(vfi_handle* h, CUtensorMap* out, const bf16* base, int cblk, int rows, int box_rows) {
  TmapKey key;
  memset(&key, 0, sizeof(key));
  key.p = base; key.d0 = cblk; key.d1 = rows; key.b0 = cblk; key.b1 = box_rows; key.rank = 2;
  auto it = tmap_cache().find(key);
  if (it != tmap_cache().end()) { *out = it->second; return 0; }
  PFN_encodeTiled enc;
  if (int rc = get_encode(h, &enc)) return rc;
  cuuint64_t dims[2] = {(cuuint64_t)cblk, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)cblk * 2};
  cuuint32_t box[2] = {(cuuint32_t)cblk, (cuuint32_t)box_rows};
  cuuint32_t es[2] = {1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, (void*)base, dims, strides, box, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle_for(cblk), CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) VFI_FAIL(h, VFI_E_CUDA, "cuTensorMapEncodeTiled(weights) failed: %d", (int)r);
  tmap_cache()[key] = *out;
  return 0;
}

int k_conv_tc(vfi_handle* h, const Layer& L, const bf16* in, const bf16* residual, void* out, int N, int Hin,
              int Win, bool out_phase, float slope, cudaStream_t s, int low_c, const LowPlanes* planes) {
  ProfScope _ps(h, PC_CONV_TC, s);
  TcParams P;
  memset(&P, 0, sizeof(P));
  const bool s2 = (L.kind == LK_CONV3 && L.stride == 2);
  P.N = N;
  P.Ho = s2 ? Hin / 2 : Hin;
  P.Wo = s2 ? Win / 2 : Win;
  P.planes = s2 ? 4 : 1;
  if (s2 && ((Hin | Win) & 1)) VFI_FAIL(h, VFI_E_ARG, "stride-2 conv needs even input dims");
  if (out_phase && ((P.Ho | P.Wo) & 1)) VFI_FAIL(h, VFI_E_ARG, "phase-plane output needs even dims");
  // tile shape: 128 pixels, minimise padded area
  static const int cand[8][2] = {{8, 16}, {16, 8}, {4, 32}, {32, 4}, {2, 64}, {64, 2}, {1, 128}, {128, 1}};
  long long best = -1;
  for (int i = 0; i < 8; i++) {
    long long a = (long long)cdiv(P.Ho, cand[i][0]) * cand[i][0] * cdiv(P.Wo, cand[i][1]) * cand[i][1];
    if (best < 0 || a < best) { best = a; P.th = cand[i][0]; P.tw = cand[i][1]; }
  }
  P.tiles_x = cdiv(P.Wo, P.tw);
  P.tiles_y = cdiv(P.Ho, P.th);
  P.num_tiles = N * P.tiles_x * P.tiles_y;
  P.nsteps = L.nsteps;
  P.cblk = L.cblk;
  P.ksub = L.cblk / 16;
  P.a_bytes = 128 * L.cblk * 2;
  int bmax = 0;
  for (int i = 0; i < L.nsteps; i++) bmax = std::max(bmax, (int)L.steps[i].b_rows);
  int b_bytes = bmax * L.cblk * 2;
  P.stage_bytes = (P.a_bytes + b_bytes + 1023) & ~1023;
  P.nstages = std::min(TC_MAX_STAGES, (200 * 1024) / P.stage_bytes);
  if (P.nstages < 2) VFI_FAIL(h, VFI_E_ARG, "tc stage too large");
  P.n_total = L.n_total;
  P.acc_stride = L.n_total <= 32 ? 32 : L.n_total <= 64 ? 64 : L.n_total <= 128 ? 128 : 256;
  P.tmem_cols = std::max(32, 2 * P.acc_stride);
  P.b_box = L.b_box;
  P.sbo16 = (uint32_t)(8 * L.cblk * 2) >> 4;
  P.layout_type = L.cblk == 64 ? 2 : L.cblk == 32 ? 4 : 6;
  P.epi_mode = L.kind == LK_DECONV4 ? 2 : (out_phase ? 1 : 0);
  P.cout = L.cout;
  P.low_c = low_c;
  P.act = L.act ? 1 : 0;
  P.slope = slope;
  P.bias = L.bias;
  P.beta = L.has_beta ? L.beta : nullptr;
  P.res = residual;
  if (P.beta && !residual) VFI_FAIL(h, VFI_E_ARG, "ResConv needs residual");
  P.out = planes ? (void*)planes->flow : out;
  P.low_mf = planes ? planes->mf : nullptr;
  P.low_f7 = planes ? planes->f7 : nullptr;
  if (planes && (L.kind != LK_DECONV4 || low_c != 16)) VFI_FAIL(h, VFI_E_ARG, "plane output is for the 13-channel lastconv");
  memcpy(P.steps, L.steps, sizeof(TcStep) * L.nsteps);

  CUtensorMap tmA, tmB;
  if (int rc = make_tmap_act(h, &tmA, in, L.cin_pad, P.Wo, P.Ho, N * P.planes, L.cblk, P.tw, P.th)) return rc;
  if (int rc = make_tmap_w(h, &tmB, L.wtc, L.cblk, L.rows_total, L.b_box)) return rc;

  size_t smem = (size_t)P.nstages * P.stage_bytes + 1024;
  static size_t smem_set = 0;
  if (smem > smem_set) {
    VFI_CUDA(h, cudaFuncSetAttribute(conv_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(208 * 1024)));
    smem_set = 208 * 1024;
  }
  int grid = std::min(P.num_tiles, h->sm_count);
  conv_tc_kernel<<<grid, TC_THREADS, smem, s>>>(tmA, tmB, P);
  VFI_LAUNCHED(h);
  return 0;
}
