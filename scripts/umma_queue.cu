// Cost of each step of the MMA-issuer loop (converged warp, elected lane issues), measured with clock64 stamps.
#include <cstdio>
#include "../p-ntfields_b200/csrc/tc_common.cuh"
using namespace pntf::tc;
__device__ __forceinline__ uint32_t mbar_test(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile("{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  return ok;
}
__global__ void __launch_bounds__(128, 1) k(long long* out, int n, int mode) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t done, rdy[4], emp[4];
  __shared__ uint32_t tmem_base_s;
  const int t = threadIdx.x, warp = t >> 5;
  if (t == 0) { mbar_init(&done, 1); for (int i = 0; i < 4; ++i) { mbar_init(&rdy[i], 1); mbar_init(&emp[i], 1); } fence_barrier_init(); }
  if (warp == 0) tmem_alloc(&tmem_base_s, 512);
  for (int i = t; i < 98304 / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  fence_proxy_async(); tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tb = tmem_base_s;
  if (t == 0) for (int i = 0; i < 4; ++i) mbar_arrive(&rdy[i]);   // all "full" barriers ready (phase 0 complete)
  __syncthreads();
  if (warp == 0) {
    const uint32_t idesc = make_idesc(128, n, false, true);
    const uint32_t a_hi = desc_hi(1024), b_hi = desc_hi(kActSBO);
    const uint32_t a_lo = desc_lo(smem_u32(smem + 65536), 16), b_lo = desc_lo(smem_u32(smem), kActLBO);
    long long ts[8][8];
    uint32_t okn = mbar_test(&rdy[0], 0);
if (mode < 3) {
#pragma unroll
    for (int it = 0; it < 4; ++it) {
      ts[it][0] = clock64();
      if (mode == 0) mbar_spin(&rdy[it], 0);
      if (mode == 1) mbar_wait(&rdy[it], 0);
      if (mode == 2) { if (!okn) mbar_spin(&rdy[it], 0); okn = mbar_test(&rdy[(it + 1) & 3], 0); }
      ts[it][1] = clock64();
      tc_fence_after();
      ts[it][2] = clock64();
      const bool el = elect_one();
      ts[it][3] = clock64();
      if (el) {
#pragma unroll
        for (int j = 0; j < 4; ++j) umma_f16_split(tb, a_lo + 2 * j, a_hi, b_lo + j * ((2 * kActSBO) >> 4), b_hi, idesc, 1u);
      }
      ts[it][4] = clock64();
      if (el) umma_commit(&emp[it]);
      ts[it][5] = clock64();
      __syncwarp();
      ts[it][6] = clock64();
    }
    } else if (mode == 3) {
#pragma unroll
      for (int it = 0; it < 4; ++it) {
        ts[it][0] = clock64();
        const bool el = elect_one();
        ts[it][1] = clock64();
        if (el) {
          mbar_spin(&rdy[it], 0);
          tc_fence_after();
#pragma unroll
          for (int j = 0; j < 4; ++j) umma_f16_split(tb, a_lo + 2 * j, a_hi, b_lo + j * ((2 * kActSBO) >> 4), b_hi, idesc, 1u);
          umma_commit(&emp[it]);
        }
        ts[it][2] = ts[it][3] = ts[it][4] = ts[it][5] = clock64();
        __syncwarp();
        ts[it][6] = clock64();
      }
    } else {
      const bool el = elect_one();
      if (el) {
#pragma unroll
        for (int it = 0; it < 4; ++it) {
          ts[it][0] = clock64();
          if (mode == 6) mbar_wait(&rdy[it], 0); else if (mode != 7) mbar_spin(&rdy[it], 0);
          ts[it][1] = clock64();
          if (mode != 8) tc_fence_after();
          ts[it][2] = ts[it][3] = clock64();
#pragma unroll
          for (int j = 0; j < 4; ++j) umma_f16_split(tb, a_lo + 2 * j, a_hi, b_lo + j * ((2 * kActSBO) >> 4), b_hi, idesc, 1u);
          ts[it][4] = clock64();
          if (mode != 5) umma_commit(&emp[it]);
          ts[it][5] = ts[it][6] = clock64();
        }
      }
      __syncwarp();
      for (int it = 0; it < 4; ++it) for (int i = 0; i < 7; ++i) ts[it][i] = __shfl_sync(0xffffffffu, ts[it][i], __ffs(__ballot_sync(0xffffffffu, el)) - 1);
    }
    if (elect_one()) umma_commit(&done);
    __syncwarp();
    mbar_spin(&done, 0);
    const long long te = clock64();
    if (t == 0) {
      for (int it = 0; it < 4; ++it)
        for (int i = 0; i < 7; ++i) out[it * 8 + i] = ts[it][i] - ts[0][0];
      out[32] = te - ts[0][0];
    }
  }
  tc_fence_before(); __syncthreads();
  if (warp == 0) tmem_dealloc(tb, 512);
}
int main() {
  long long* d; cudaMalloc(&d, 40 * 8);
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 100352);
  for (int mode : {4, 5, 6, 7, 8}) {
    const int n = 128;
    k<<<1, 128, 100352>>>(d, n, mode);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[40]; cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
    printf("mode %d N=%d %s (start, polled, fenced, elected, 4 UMMAs issued, committed, syncwarp)\n", mode, n, cudaGetErrorString(e));
    for (int it = 0; it < 4; ++it) { for (int i = 0; i < 7; ++i) printf(" %5lld", h[it * 8 + i]); printf("\n"); }
    printf(" all done %lld\n", h[32]);
  }
}
