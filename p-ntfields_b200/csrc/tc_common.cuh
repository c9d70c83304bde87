// sm_100a building blocks of the tensor-core precision mode: mbarrier, bulk async copies (TMA 1-D),
// tcgen05 (TMEM allocation, UMMA issue/commit, TMEM load/store) and the shared-memory operand
// layouts.  Inline PTX only; no CUTLASS.
//
// Operand layouts (fp16, SWIZZLE_128B, "atom" = 8 lines of 128 bytes, 1024-byte aligned):
//   * weight stage  A[128 out-features x 64 in-features], K-major: feature m lives in atom m/8, line m%8;
//     the 16-byte chunk c (= 8 consecutive k) of a line is stored at chunk position c ^ (m%8).
//   * activation tile B[K features x R rows], MN-major: feature k lives in atom (k/8, r/64), line k%8;
//     the 8 consecutive rows of chunk c are stored at chunk position c ^ (k%8).
//   Both are the same physical pattern (one 128-byte line per feature, chunk index XOR line index),
//   so a tile written as "B, MN-major" can also be consumed as a K-major operand whose K is the row
//   axis (the dW contraction).
#pragma once
#include <cstdint>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

namespace pntf {
namespace tc {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

// ---------------------------------------------------------------- mbarrier
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// Bounded wait: a protocol bug traps (launch error) instead of hanging the GPU.  try_wait itself
// suspends the thread for a hardware-defined interval, so the loop is not a hot spin.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  uint32_t spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    if (++spins > (1u << 28)) __trap();
  }
}
__device__ __forceinline__ uint32_t ld_acquire_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.cta.shared.u32 %0, [%1];" : "=r"(v) : "r"(smem_u32(p)) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_u32(uint32_t* p, uint32_t v) {
  asm volatile("st.release.cta.shared.u32 [%0], %1;" ::"r"(smem_u32(p)), "r"(v) : "memory");
}
// Low-latency variant for single issuer threads: polls with test_wait (no suspend hint).
__device__ __forceinline__ void mbar_spin(uint64_t* bar, uint32_t parity) {
  uint32_t spins = 0;
  while (true) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    if (ok) return;
    if (++spins > (1u << 30)) __trap();
  }
}
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
// generic-proxy writes to shared memory -> visible to the async proxy (UMMA operand reads, bulk stores)
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ---------------------------------------------------------------- bulk copies (TMA, 1-D)
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(smem_dst)),
               "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void bulk_s2g(void* gmem_dst, const void* smem_src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gmem_dst), "r"(smem_u32(smem_src)),
               "r"(bytes)
               : "memory");
}
// L2 eviction priority of the stash traffic (-DPNTF_L2_HINTS=0 turns it off).  Every stash line is written once and read
// once or twice, gigabytes later: it can never hit, but at normal priority it pushes the weight images, biases and
// coordinates out of L2.  The stash stores of the training forward and the stash loads / Ybar stores of the reverse sweep
// carry evict_first: same box A/B 15.00 -> 15.42 and 15.85 -> 16.2 M pairs/s (reverse sweep -1.6 ms, forward -0.5 ms per
// step; profiles/r02q_variants.txt).  The weight-gradient kernel keeps normal priority: the X tiles of the 256-feature
// layers are read by the two work units (output halves) of the layer at about the same time and the second read must hit
// (with evict_first on its loads it lost 0.4 ms per step); evict_last on the weight stages adds nothing on top.
#ifndef PNTF_L2_HINTS
#define PNTF_L2_HINTS 1
#endif
__device__ __forceinline__ uint64_t l2_evict_first() {
  uint64_t p;
  asm("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ void stash_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
#if PNTF_L2_HINTS
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(
                   smem_u32(smem_dst)),
               "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar)), "l"(l2_evict_first())
               : "memory");
#else
  bulk_g2s(smem_dst, gmem_src, bytes, bar);
#endif
}
__device__ __forceinline__ void stash_s2g(void* gmem_dst, const void* smem_src, uint32_t bytes) {
#if PNTF_L2_HINTS
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group.L2::cache_hint [%0], [%1], %2, %3;" ::"l"(gmem_dst),
               "r"(smem_u32(smem_src)), "r"(bytes), "l"(l2_evict_first())
               : "memory");
#else
  bulk_s2g(gmem_dst, smem_src, bytes);
#endif
}
// 4-byte asynchronous copy global -> shared (LDGSTS: no register holds the value while it is in flight); a thread sees its
// own copies after cp_async_wait_all()
__device__ __forceinline__ void cp_async4(void* smem_dst, const void* gmem_src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(smem_dst)), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }
// asynchronous prefetch of a global range into L2 (no destination, no completion to wait for)
__device__ __forceinline__ void bulk_prefetch_l2(const void* gmem_src, uint32_t bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(gmem_src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// all committed bulk stores have finished READING shared memory
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
// all but the N most recent bulk-store groups have finished reading shared memory
template <int N>
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
// all but the N most recent bulk-store groups have completed (their writes are performed)
template <int N>
__device__ __forceinline__ void bulk_wait() { asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void bulk_wait0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

// ---------------------------------------------------------------- tcgen05 / TMEM
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_result, uint32_t ncols) {   // one full warp
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_result)), "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {        // the same warp
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// one lane of a converged warp (the pattern ptxas recognises as single-thread-uniform issue)
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "elect.sync _|p, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(pred));
  return pred != 0;
}

// D[tmem] (+)= A[smem] * B[smem], fp16 operands, fp32 accumulate; one thread issues for the CTA.
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on `bar` once every previously issued UMMA of this thread has completed
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// Shared-memory operand descriptor, SWIZZLE_128B (sm_100 "version 1").  lbo/sbo in bytes.
__device__ __forceinline__ uint64_t make_desc(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= uint64_t((smem_addr >> 4) & 0x3FFF);
  d |= uint64_t((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= uint64_t((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= uint64_t(1) << 46;    // descriptor version (Blackwell)
  d |= uint64_t(2) << 61;    // SWIZZLE_128B
  return d;
}
// The same descriptor split in two 32-bit words: the address field lives in lo[0,14), so stepping the
// operand by n bytes is lo + (n >> 4) (no carry into the LBO field for shared-memory addresses).
__device__ __forceinline__ uint32_t desc_lo(uint32_t smem_addr, uint32_t lbo_bytes) {
  return ((smem_addr >> 4) & 0x3FFF) | (((lbo_bytes >> 4) & 0x3FFF) << 16);
}
__host__ __device__ constexpr uint32_t desc_hi(uint32_t sbo_bytes) {
  return ((sbo_bytes >> 4) & 0x3FFF) | (1u << 14) | (2u << 29);
}
__device__ __forceinline__ void umma_f16_split(uint32_t tmem_d, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo, uint32_t b_hi,
                                               uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\t"
      "setp.ne.b32 p, %6, 0;\n\t"
      "mov.b64 da, {%1, %2};\n\t"
      "mov.b64 db, {%3, %4};\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %5, p;\n\t}" ::"r"(tmem_d),
      "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate)
      : "memory");
}
// Instruction descriptor, kind::f16: fp16 x fp16 -> fp32.  a_mn / b_mn: operand is MN-major.
__host__ __device__ constexpr uint32_t make_idesc(int M, int N, bool a_mn, bool b_mn) {
  return (1u << 4) | (uint32_t(a_mn) << 15) | (uint32_t(b_mn) << 16) | (uint32_t(N >> 3) << 17) | (uint32_t(M >> 4) << 24);
}

// TMEM address: lane in bits [16,32), column in bits [0,16)
__device__ __forceinline__ uint32_t tmem_addr(uint32_t base, uint32_t lane, uint32_t col) { return base + (lane << 16) + col; }

// 16 consecutive fp32 columns of this thread's lane (warp w reads lanes 32*(w%4) .. +31)
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16]) {
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const float (&v)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};" ::"r"(taddr),
      "r"(__float_as_uint(v[0])), "r"(__float_as_uint(v[1])), "r"(__float_as_uint(v[2])), "r"(__float_as_uint(v[3])),
      "r"(__float_as_uint(v[4])), "r"(__float_as_uint(v[5])), "r"(__float_as_uint(v[6])), "r"(__float_as_uint(v[7])),
      "r"(__float_as_uint(v[8])), "r"(__float_as_uint(v[9])), "r"(__float_as_uint(v[10])), "r"(__float_as_uint(v[11])),
      "r"(__float_as_uint(v[12])), "r"(__float_as_uint(v[13])), "r"(__float_as_uint(v[14])), "r"(__float_as_uint(v[15]))
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// NC consecutive fp32 columns of this thread's lane (NC = 8, 16, 32).  The wait lists the destination
// registers as in/out operands so that no use of them can be scheduled before it.
template <int NC> __device__ __forceinline__ void tmem_ld(uint32_t taddr, float* v);
template <int NC> __device__ __forceinline__ void tmem_st(uint32_t taddr, const float* v);   // needs tmem_st_wait()
template <> __device__ __forceinline__ void tmem_ld<8>(uint32_t taddr, float* v) {
  uint32_t r[8];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(r[0]),"=r"(r[1]),"=r"(r[2]),"=r"(r[3]),"=r"(r[4]),"=r"(r[5]),"=r"(r[6]),"=r"(r[7])
               : "r"(taddr)
               : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" : "+r"(r[0]),"+r"(r[1]),"+r"(r[2]),"+r"(r[3]),"+r"(r[4]),"+r"(r[5]),"+r"(r[6]),"+r"(r[7]) : : "memory");
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(r[i]);
}
template <> __device__ __forceinline__ void tmem_st<8>(uint32_t taddr, const float* v) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(taddr), "r"(__float_as_uint(v[0])),"r"(__float_as_uint(v[1])),"r"(__float_as_uint(v[2])),"r"(__float_as_uint(v[3])),"r"(__float_as_uint(v[4])),"r"(__float_as_uint(v[5])),"r"(__float_as_uint(v[6])),"r"(__float_as_uint(v[7]))
               : "memory");
}
template <> __device__ __forceinline__ void tmem_ld<16>(uint32_t taddr, float* v) {
  uint32_t r[16];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
               : "=r"(r[0]),"=r"(r[1]),"=r"(r[2]),"=r"(r[3]),"=r"(r[4]),"=r"(r[5]),"=r"(r[6]),"=r"(r[7]),"=r"(r[8]),"=r"(r[9]),"=r"(r[10]),"=r"(r[11]),"=r"(r[12]),"=r"(r[13]),"=r"(r[14]),"=r"(r[15])
               : "r"(taddr)
               : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" : "+r"(r[0]),"+r"(r[1]),"+r"(r[2]),"+r"(r[3]),"+r"(r[4]),"+r"(r[5]),"+r"(r[6]),"+r"(r[7]),"+r"(r[8]),"+r"(r[9]),"+r"(r[10]),"+r"(r[11]),"+r"(r[12]),"+r"(r[13]),"+r"(r[14]),"+r"(r[15]) : : "memory");
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}
template <> __device__ __forceinline__ void tmem_st<16>(uint32_t taddr, const float* v) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};" ::"r"(taddr), "r"(__float_as_uint(v[0])),"r"(__float_as_uint(v[1])),"r"(__float_as_uint(v[2])),"r"(__float_as_uint(v[3])),"r"(__float_as_uint(v[4])),"r"(__float_as_uint(v[5])),"r"(__float_as_uint(v[6])),"r"(__float_as_uint(v[7])),"r"(__float_as_uint(v[8])),"r"(__float_as_uint(v[9])),"r"(__float_as_uint(v[10])),"r"(__float_as_uint(v[11])),"r"(__float_as_uint(v[12])),"r"(__float_as_uint(v[13])),"r"(__float_as_uint(v[14])),"r"(__float_as_uint(v[15]))
               : "memory");
}
template <> __device__ __forceinline__ void tmem_ld<32>(uint32_t taddr, float* v) {
  uint32_t r[32];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
               : "=r"(r[0]),"=r"(r[1]),"=r"(r[2]),"=r"(r[3]),"=r"(r[4]),"=r"(r[5]),"=r"(r[6]),"=r"(r[7]),"=r"(r[8]),"=r"(r[9]),"=r"(r[10]),"=r"(r[11]),"=r"(r[12]),"=r"(r[13]),"=r"(r[14]),"=r"(r[15]),"=r"(r[16]),"=r"(r[17]),"=r"(r[18]),"=r"(r[19]),"=r"(r[20]),"=r"(r[21]),"=r"(r[22]),"=r"(r[23]),"=r"(r[24]),"=r"(r[25]),"=r"(r[26]),"=r"(r[27]),"=r"(r[28]),"=r"(r[29]),"=r"(r[30]),"=r"(r[31])
               : "r"(taddr)
               : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" : "+r"(r[0]),"+r"(r[1]),"+r"(r[2]),"+r"(r[3]),"+r"(r[4]),"+r"(r[5]),"+r"(r[6]),"+r"(r[7]),"+r"(r[8]),"+r"(r[9]),"+r"(r[10]),"+r"(r[11]),"+r"(r[12]),"+r"(r[13]),"+r"(r[14]),"+r"(r[15]),"+r"(r[16]),"+r"(r[17]),"+r"(r[18]),"+r"(r[19]),"+r"(r[20]),"+r"(r[21]),"+r"(r[22]),"+r"(r[23]),"+r"(r[24]),"+r"(r[25]),"+r"(r[26]),"+r"(r[27]),"+r"(r[28]),"+r"(r[29]),"+r"(r[30]),"+r"(r[31]) : : "memory");
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}
template <> __device__ __forceinline__ void tmem_st<32>(uint32_t taddr, const float* v) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};" ::"r"(taddr), "r"(__float_as_uint(v[0])),"r"(__float_as_uint(v[1])),"r"(__float_as_uint(v[2])),"r"(__float_as_uint(v[3])),"r"(__float_as_uint(v[4])),"r"(__float_as_uint(v[5])),"r"(__float_as_uint(v[6])),"r"(__float_as_uint(v[7])),"r"(__float_as_uint(v[8])),"r"(__float_as_uint(v[9])),"r"(__float_as_uint(v[10])),"r"(__float_as_uint(v[11])),"r"(__float_as_uint(v[12])),"r"(__float_as_uint(v[13])),"r"(__float_as_uint(v[14])),"r"(__float_as_uint(v[15])),"r"(__float_as_uint(v[16])),"r"(__float_as_uint(v[17])),"r"(__float_as_uint(v[18])),"r"(__float_as_uint(v[19])),"r"(__float_as_uint(v[20])),"r"(__float_as_uint(v[21])),"r"(__float_as_uint(v[22])),"r"(__float_as_uint(v[23])),"r"(__float_as_uint(v[24])),"r"(__float_as_uint(v[25])),"r"(__float_as_uint(v[26])),"r"(__float_as_uint(v[27])),"r"(__float_as_uint(v[28])),"r"(__float_as_uint(v[29])),"r"(__float_as_uint(v[30])),"r"(__float_as_uint(v[31]))
               : "memory");
}
// fire-and-forget fp32 add to a shared-memory word (shared-window address)
__device__ __forceinline__ void red_shared_add(uint32_t addr, float v) {
  asm volatile("red.shared.add.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}
// barrier among a subset of the CTA's warps (id 1..15)
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory"); }

// ---------------------------------------------------------------- operand tile geometry
constexpr int kStageBytes = 16384;          // one weight stage: 128 x 64 fp16
constexpr int kAtomBytes = 1024;

// byte offset of element (m, k) inside a K-major weight stage [128 x 64]
__host__ __device__ constexpr int wstage_off(int m, int k) {
  return (m >> 3) * kAtomBytes + (m & 7) * 128 + ((((k >> 3) ^ (m & 7)) & 7) << 4) + (k & 7) * 2;
}
// Activation tile [K features x 128 rows]: atoms ordered (feature group kg = k/8, row block nb = r/64).
// byte offset of the 16-byte chunk holding rows [8c, 8c+8) of row block nb for feature k
__host__ __device__ constexpr int act_chunk_off(int k, int nb, int c) {
  return ((k >> 3) * 2 + nb) * kAtomBytes + (k & 7) * 128 + (((c ^ (k & 7)) & 7) << 4);
}
constexpr int kActLBO = kAtomBytes;         // between the two 64-row blocks of a feature group
constexpr int kActSBO = 2 * kAtomBytes;     // between consecutive groups of 8 features

// two floats -> fp16x2 (a in the low half), round to nearest, SATURATING: a magnitude beyond 65504 becomes +-65504
// instead of inf, so that one out-of-range element cannot turn a whole tile into NaNs in the next product
__device__ __forceinline__ uint32_t pack_h2(float a, float b) {
  uint32_t r;
  asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a));
  return r;
}
// write 8 consecutive rows (one chunk) of one feature
__device__ __forceinline__ void st_chunk(uint8_t* tile, int k, int nb, int c, const float* v) {
  uint4 q;
  q.x = pack_h2(v[0], v[1]); q.y = pack_h2(v[2], v[3]); q.z = pack_h2(v[4], v[5]); q.w = pack_h2(v[6], v[7]);
  *reinterpret_cast<uint4*>(tile + act_chunk_off(k, nb, c)) = q;
}

// same, addressed by a shared-window byte address with the per-thread part of the swizzle hoisted:
// line = address of the 128-byte line of feature k in row block nb, kx = k & 7, c = chunk 0..7
__device__ __forceinline__ uint32_t act_line_addr(uint32_t tile_addr, int k, int nb) {
  return tile_addr + uint32_t(((k >> 3) * 2 + nb) * kAtomBytes + (k & 7) * 128);
}
__device__ __forceinline__ void sts_chunk(uint32_t line, int kx, int c, const float* v) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(line + uint32_t(((c ^ kx) & 7) << 4)), "r"(pack_h2(v[0], v[1])),
               "r"(pack_h2(v[2], v[3])), "r"(pack_h2(v[4], v[5])), "r"(pack_h2(v[6], v[7]))
               : "memory");
}
__device__ __forceinline__ void lds_chunk(uint32_t line, int kx, int c, float* v) {
  uint32_t q[4];
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(q[0]), "=r"(q[1]), "=r"(q[2]), "=r"(q[3])
               : "r"(line + uint32_t(((c ^ kx) & 7) << 4))
               : "memory");
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 t = __half22float2(*reinterpret_cast<const __half2*>(&q[i]));
    v[2 * i] = t.x;
    v[2 * i + 1] = t.y;
  }
}

}  // namespace tc
}  // namespace pntf
