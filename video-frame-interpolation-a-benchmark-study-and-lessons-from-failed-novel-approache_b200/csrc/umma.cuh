// umma.cuh -- the PTX wrappers (mbarrier, cp.async, tcgen05 alloc / mma / commit / ld, no-swizzle smem descriptors) shared
// by the resident-weight implicit-GEMM kernels (conv_s2.cu and the fused gather kernels of elementwise.cu and conv_v2.cu).  sm_100a only.
#pragma once
#include "common.cuh"

namespace umma {
__device__ __forceinline__ uint32_t s_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void bar_init(uint64_t* b, uint32_t cnt) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s_u32(b)), "r"(cnt) : "memory");
}
__device__ __forceinline__ void bar_wait(uint64_t* b, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred P1;\n\t"
      "UM_WAIT:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t"
      "@P1 bra UM_DONE;\n\t"
      "bra UM_WAIT;\n\t"
      "UM_DONE:\n\t}" ::"r"(s_u32(b)), "r"(parity)
      : "memory");
}
__device__ __forceinline__ void bar_arrive(uint64_t* b) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(s_u32(b)) : "memory");
}
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, uint32_t src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void tc_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_alloc(uint32_t* slot, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s_u32(slot)), "r"(cols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t base, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(base), "r"(cols) : "memory");
}
__device__ __forceinline__ void commit_to(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(s_u32(bar)) : "memory");
}
// issued from a converged warp: `leader` is the elect.sync result, operands are warp-uniform
__device__ __forceinline__ void mma_bf16_elect(uint32_t leader, uint32_t d_tmem, uint64_t adesc, uint64_t bdesc,
                                               uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p, pe;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "setp.ne.b32 pe, %5, 0;\n\t"
      "@pe tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate), "r"(leader)
      : "memory");
}
__device__ __forceinline__ uint32_t elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred pe;\n\telect.sync _|pe, 0xffffffff;\n\tselp.u32 %0, 1, 0, pe;\n\t}" : "=r"(pred));
  return pred;
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
// kind::f16 instruction descriptor: F32 accumulate, BF16 x BF16, K-major A and B, M = 128, N = n
__host__ __device__ constexpr uint32_t idesc_bf16_m128(int n) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | (8u << 24);
}
__device__ __forceinline__ uint32_t pk_bf16(float a, float b) {
  __nv_bfloat162 t = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&t);
}
static inline uint16_t f2bf_host(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffff) > 0x7f800000) return 0x7fc0;
  u += 0x7fff + ((u >> 16) & 1);
  return (uint16_t)(u >> 16);
}
// chunk stride (16-byte units) >= npix such that the 8 lanes of a 16-byte store phase hit 8 distinct bank groups
constexpr int chunk_stride(int npix, int nchunk) {
  int c = npix;
  if (nchunk >= 6) { while (c % 2 == 0) c++; }
  else if (nchunk == 4) { while (c % 8 != 2) c++; }
  else if (nchunk == 2) { while (c % 8 != 4) c++; }
  return c;
}
// A persistent CTA's walk over its tiles (tile0, tile0 + step, ...) WITHOUT per-tile integer divisions: the linear index is
// decoded to (tx, ty, n) once and the mixed-radix coordinates then advance by the decomposed step.  The three runtime
// div / mod of the decode cost ~100-150 cycles each, dependent, in EVERY role of a warp-specialised kernel; with the
// it % nstages of the stage ring they were ~860 cycles per tile on the epilogue warps' critical path (tools/v2_trace.py).
struct TileWalk {
  int tile, tx, ty, n, step, sx, sy, sn, tiles_x, tiles_y;
  __device__ __forceinline__ TileWalk(int tile0, int step_, int tiles_x_, int tiles_y_)
      : tile(tile0), step(step_), tiles_x(tiles_x_), tiles_y(tiles_y_) {
    const int per = tiles_x * tiles_y;
    tx = tile0 % tiles_x; ty = (tile0 / tiles_x) % tiles_y; n = tile0 / per;
    sx = step_ % tiles_x; sy = (step_ / tiles_x) % tiles_y; sn = step_ / per;
  }
  __device__ __forceinline__ void next() {
    tile += step; tx += sx; ty += sy; n += sn;
    if (tx >= tiles_x) { tx -= tiles_x; ty++; }
    if (ty >= tiles_y) { ty -= tiles_y; n++; }
  }
};
// stage index and mbarrier phase parity of iteration it = 0, 1, ... of an nst-deep ring: st = it % nst, ph = (it / nst) & 1
struct StageWalk {
  int st, nst; uint32_t ph;
  __device__ __forceinline__ explicit StageWalk(int nst_, int start = 0) : st(start % nst_), nst(nst_), ph((uint32_t)((start / nst_) & 1)) {}
  __device__ __forceinline__ void next() { if (++st == nst) { st = 0; ph ^= 1u; } }
};
}  // namespace umma
