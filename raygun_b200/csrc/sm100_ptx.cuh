// Thin inline-PTX wrappers for the sm_100a features the implicit-GEMM kernels use:
// mbarrier, TMA (cp.async.bulk.tensor), tcgen05 (alloc / mma / commit / ld / fences).
#pragma once
#include <cuda.h>
#include <stdint.h>

namespace rg {
namespace ptx {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ---- mbarrier ----------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
// Bounded wait: a protocol bug must trap (error surfaced to the host), never hang the GPU.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  long long t0 = clock64();
  while (!mbar_try_wait(bar, parity)) {
    if (clock64() - t0 > 4000000000LL) {  // ~2 s at 2 GHz
      printf("raygun_b200: mbarrier wait timed out (block %d thread %d bar %u parity %u)\n", (int)blockIdx.x,
             (int)threadIdx.x, bar, parity);
      __trap();
    }
  }
}

// Long waits of roles that are idle most of the time (epilogue, producer): back off so the spinning warp does not
// take issue slots from the MMA-issuing warp that shares its scheduler.
__device__ __forceinline__ void mbar_wait_backoff(uint32_t bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  long long t0 = clock64();
  while (!mbar_try_wait(bar, parity)) {
    __nanosleep(64);
    if (clock64() - t0 > 4000000000LL) {
      printf("raygun_b200: mbarrier wait timed out (block %d thread %d bar %u parity %u)\n", (int)blockIdx.x,
             (int)threadIdx.x, bar, parity);
      __trap();
    }
  }
}
// One lane of a fully active warp (convergent callers only).
__device__ __forceinline__ bool elect_one_sync() {
  uint32_t pred;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "elect.sync _|p, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(pred));
  return pred != 0;
}

// ---- TMA ---------------------------------------------------------------------------------------
__device__ __forceinline__ void prefetch_tmap(const CUtensorMap* m) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(m) : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* m, uint32_t bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(dst),
      "l"(m), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_load_5d(uint32_t dst, const CUtensorMap* m, uint32_t bar, int c0, int c1, int c2,
                                            int c3, int c4) {
  asm volatile(
      "cp.async.bulk.tensor.5d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6, %7}], "
      "[%2];" ::"r"(dst),
      "l"(m), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
      : "memory");
}
__device__ __forceinline__ void tma_store_5d(const CUtensorMap* m, uint32_t src, int c0, int c1, int c2, int c3, int c4) {
  asm volatile("cp.async.bulk.tensor.5d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5, %6}], [%1];" ::"l"(m),
               "r"(src), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
               : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void tma_store_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N> __device__ __forceinline__ void tma_store_wait_all() {
  asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ---- tcgen05 -----------------------------------------------------------------------------------
template <int COLS> __device__ __forceinline__ void tmem_alloc(uint32_t smem_dst) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_dst), "n"(COLS) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
template <int COLS> __device__ __forceinline__ void tmem_dealloc(uint32_t taddr) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(COLS) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// D[tmem] (+)= A[smem desc] * B[smem desc], bf16 x bf16 -> fp32
__device__ __forceinline__ void umma_bf16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// mbarrier arrives when all previously issued MMAs of this thread have completed
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}

// Predicated forms for a convergent issuing warp: every lane executes the instruction slot, only lanes with
// `issue != 0` (the elected leader) issue -- no divergent branch, hence no BSSY/BSYNC reconvergence per MMA group.
__device__ __forceinline__ void umma_bf16_pred(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                               uint32_t accumulate, uint32_t issue) {
  asm volatile(
      "{\n\t.reg .pred p, q;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "setp.ne.b32 q, %5, 0;\n\t"
      "@q tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate), "r"(issue)
      : "memory");
}
// the four K = 16 steps of one 64-channel chunk: descriptor low words advance by 32 B (>> 4 = 2) per step; always accumulates
__device__ __forceinline__ void umma_bf16_x4_pred(uint32_t d_tmem, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo, uint32_t b_hi,
                                                  uint32_t idesc, uint32_t issue) {
  asm volatile(
      "{\n\t.reg .pred q, one;\n\t.reg .b64 a, b;\n\t.reg .b32 ta, tb;\n\t"
      "setp.ne.b32 q, %6, 0;\n\t"
      "setp.eq.b32 one, 0, 0;\n\t"
      "mov.b64 a, {%1, %2};\n\t"
      "mov.b64 b, {%3, %4};\n\t"
      "@q tcgen05.mma.cta_group::1.kind::f16 [%0], a, b, %5, one;\n\t"
      "add.u32 ta, %1, 2;\n\t add.u32 tb, %3, 2;\n\t mov.b64 a, {ta, %2};\n\t mov.b64 b, {tb, %4};\n\t"
      "@q tcgen05.mma.cta_group::1.kind::f16 [%0], a, b, %5, one;\n\t"
      "add.u32 ta, %1, 4;\n\t add.u32 tb, %3, 4;\n\t mov.b64 a, {ta, %2};\n\t mov.b64 b, {tb, %4};\n\t"
      "@q tcgen05.mma.cta_group::1.kind::f16 [%0], a, b, %5, one;\n\t"
      "add.u32 ta, %1, 6;\n\t add.u32 tb, %3, 6;\n\t mov.b64 a, {ta, %2};\n\t mov.b64 b, {tb, %4};\n\t"
      "@q tcgen05.mma.cta_group::1.kind::f16 [%0], a, b, %5, one;\n\t}" ::"r"(d_tmem),
      "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(issue)
      : "memory");
}
__device__ __forceinline__ void umma_bf16_x4(uint32_t d_tmem, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo, uint32_t b_hi,
                                             uint32_t idesc) {
  asm volatile(
      "{\n\t.reg .pred one;\n\t.reg .b64 a, b;\n\t.reg .b32 ta, tb;\n\t"
      "setp.eq.b32 one, 0, 0;\n\t"
      "mov.b64 a, {%1, %2};\n\t"
      "mov.b64 b, {%3, %4};\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], a, b, %5, one;\n\t"
      "add.u32 ta, %1, 2;\n\t add.u32 tb, %3, 2;\n\t mov.b64 a, {ta, %2};\n\t mov.b64 b, {tb, %4};\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], a, b, %5, one;\n\t"
      "add.u32 ta, %1, 4;\n\t add.u32 tb, %3, 4;\n\t mov.b64 a, {ta, %2};\n\t mov.b64 b, {tb, %4};\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], a, b, %5, one;\n\t"
      "add.u32 ta, %1, 6;\n\t add.u32 tb, %3, 6;\n\t mov.b64 a, {ta, %2};\n\t mov.b64 b, {tb, %4};\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], a, b, %5, one;\n\t}" ::"r"(d_tmem),
      "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc)
      : "memory");
}
// eight K = 16 steps (one 128-voxel chunk of a weight-gradient GEMM): descriptor low words advance by a_step / b_step;
// the first step accumulates iff acc_first, the others always
__device__ __forceinline__ void umma_bf16_x8(uint32_t d_tmem, uint32_t a_lo, uint32_t a_step, uint32_t a_hi, uint32_t b_lo,
                                             uint32_t b_step, uint32_t b_hi, uint32_t idesc, uint32_t acc_first) {
  asm volatile(
      "{\n\t.reg .pred pa, one;\n\t.reg .b64 a, b;\n\t.reg .b32 ta, tb;\n\t"
      "setp.ne.b32 pa, %8, 0;\n\t"
      "setp.eq.b32 one, 0, 0;\n\t"
      "mov.b32 ta, %1;\n\t mov.b32 tb, %4;\n\t mov.b64 a, {ta, %3};\n\t mov.b64 b, {tb, %6};\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], a, b, %7, pa;\n\t"
      "add.u32 ta, ta, %2;\n\t add.u32 tb, tb, %5;\n\t mov.b64 a, {ta, %3};\n\t mov.b64 b, {tb, %6};\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], a, b, %7, one;\n\t"
      "add.u32 ta, ta, %2;\n\t add.u32 tb, tb, %5;\n\t mov.b64 a, {ta, %3};\n\t mov.b64 b, {tb, %6};\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], a, b, %7, one;\n\t"
      "add.u32 ta, ta, %2;\n\t add.u32 tb, tb, %5;\n\t mov.b64 a, {ta, %3};\n\t mov.b64 b, {tb, %6};\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], a, b, %7, one;\n\t"
      "add.u32 ta, ta, %2;\n\t add.u32 tb, tb, %5;\n\t mov.b64 a, {ta, %3};\n\t mov.b64 b, {tb, %6};\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], a, b, %7, one;\n\t"
      "add.u32 ta, ta, %2;\n\t add.u32 tb, tb, %5;\n\t mov.b64 a, {ta, %3};\n\t mov.b64 b, {tb, %6};\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], a, b, %7, one;\n\t"
      "add.u32 ta, ta, %2;\n\t add.u32 tb, tb, %5;\n\t mov.b64 a, {ta, %3};\n\t mov.b64 b, {tb, %6};\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], a, b, %7, one;\n\t"
      "add.u32 ta, ta, %2;\n\t add.u32 tb, tb, %5;\n\t mov.b64 a, {ta, %3};\n\t mov.b64 b, {tb, %6};\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], a, b, %7, one;\n\t"
      "}" ::"r"(d_tmem),
      "r"(a_lo), "r"(a_step), "r"(a_hi), "r"(b_lo), "r"(b_step), "r"(b_hi), "r"(idesc), "r"(acc_first)
      : "memory");
}
// a warp-uniform value routed through REDUX so that it lives in a uniform register (and what is derived from it
// stays in the uniform datapath: tcgen05.mma takes all its operands from there)
__device__ __forceinline__ uint32_t uniform_u32(uint32_t v) { return __reduce_or_sync(0xffffffffu, v); }
__device__ __forceinline__ void umma_commit_pred(uint32_t bar, uint32_t issue) {
  asm volatile(
      "{\n\t.reg .pred q;\n\t"
      "setp.ne.b32 q, %1, 0;\n\t"
      "@q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}" ::"r"(bar),
      "r"(issue)
      : "memory");
}

// 32 lanes x 32 columns of fp32: thread i of the warp receives columns [c, c+32) of lane (base_lane+i)
__device__ __forceinline__ void tmem_ld_32x32(uint32_t taddr, uint32_t* v) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
// Store the same 32-bit value to 32 lanes x 32 columns (used to re-zero drained accumulator blocks).
__device__ __forceinline__ void tmem_st_32x32_fill(uint32_t taddr, uint32_t z) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, "
      "%1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1};" ::"r"(taddr),
      "r"(z)
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

__device__ __forceinline__ void named_bar_sync(int id, int threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

// ---- descriptors ---------------------------------------------------------------------------------
// Shared-memory matrix descriptor (cute::UMMA::SmemDescriptor layout, SWIZZLE_128B).
//   [0,14) start>>4 | [16,30) LBO>>4 | [32,46) SBO>>4 | [46,48) version=1 | [61,64) layout (2 = SW128)
__host__ __device__ __forceinline__ uint64_t smem_desc_sw128(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
// Instruction descriptor, kind::f16: bf16 A/B, fp32 D.  a_major/b_major: 0 = K-major, 1 = MN-major.
__host__ __device__ __forceinline__ constexpr uint32_t idesc_bf16(int M, int N, int a_major, int b_major) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)a_major << 15) | ((uint32_t)b_major << 16) |
         ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

}  // namespace ptx
}  // namespace rg
