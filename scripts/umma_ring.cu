// The forward kernel's MMA-issuer loop in isolation (compile-time ring of 5 x 16 KB weight stages fed by one producer
// thread, 4 UMMAs N=128 per stage, poll + commit per stage), with W idle warps blocked in mbarrier.try_wait on an
// accumulator barrier, as the 16 epilogue warps are during an MMA phase.  Reports clk per UMMA.
#include <cstdio>
#include <vector>
#include "../p-ntfields_b200/csrc/tc_common.cuh"
using namespace pntf::tc;
constexpr int S = 5;
__global__ void __launch_bounds__(704, 1) k(const uint8_t* img, long long* out, int nst, int waiters, int wait_mode, int issuers) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* act = smem; uint8_t* ring = smem + 65536;
  __shared__ uint64_t full[S], empty[S], done;
  __shared__ uint32_t tmem_base_s;
  const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
  if (t == 0) { for (int i = 0; i < S; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); } mbar_init(&done, 2); fence_barrier_init(); }
  if (warp == 0) tmem_alloc(&tmem_base_s, 512);
  for (int i = t; i < (65536 + S * kStageBytes) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  fence_proxy_async(); tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tb = tmem_base_s;
  if (warp == 1) {
    if (lane == 0)
      for (int it = 0; it < nst; ++it) {
        const uint32_t slot = it % S, ph = (it / S) & 1;
        mbar_wait(&empty[slot], ph ^ 1);
        mbar_arrive_expect_tx(&full[slot], kStageBytes);
        bulk_g2s(ring + slot * kStageBytes, img + size_t(it % 68) * kStageBytes, kStageBytes, &full[slot]);
      }
  } else if (warp == 0 || (warp == 2 && issuers == 2)) {
    const int me = warp == 0 ? 0 : 1;
    constexpr uint32_t idesc = make_idesc(128, 128, false, true);
    constexpr uint32_t a_hi = desc_hi(1024), b_hi = desc_hi(kActSBO);
    const uint32_t ring_addr = smem_u32(ring), act_addr = smem_u32(act);
    long long t0 = 0;
    for (int it = (issuers == 2 ? me : 0); it < nst; it += issuers) {
      const uint32_t slot = it % S, ph = (it / S) & 1;
      mbar_spin(&full[slot], ph);
      tc_fence_after();
      if (it == 2 * S || it == 2 * S + 1) t0 = clock64();
      if (elect_one()) {
        const uint32_t a_lo = desc_lo(ring_addr + slot * kStageBytes, 16);
        const uint32_t b_lo = desc_lo(act_addr + (it & 3) * 8 * kActSBO, kActLBO);
#pragma unroll
        for (int j = 0; j < 4; ++j) umma_f16_split(tb, a_lo + 2 * j, a_hi, b_lo + j * ((2 * kActSBO) >> 4), b_hi, idesc, 1u);
        umma_commit(&empty[slot]);
      }
      __syncwarp();
    }
    if (elect_one()) { umma_commit(&done); if (issuers == 1) mbar_arrive(&done); }
    __syncwarp();
    mbar_wait(&done, 0);
    if (lane == 0 && me == 0) out[blockIdx.x] = (clock64() - t0) * 100 / ((nst - 2 * S) * 4);
  } else if (warp >= 3 && warp < 3 + waiters) {
    if (wait_mode == 0) mbar_wait(&done, 0);                       // every lane polls (the kernels' form)
    else if (wait_mode == 1) { if (lane == 0) mbar_wait(&done, 0); __syncwarp(); }
    else { while (!mbar_try_wait(&done, 0)) __nanosleep(200); }    // explicit back-off between polls
  }
  tc_fence_before(); __syncthreads();
  if (warp == 0) tmem_dealloc(tb, 512);
}
int main() {
  uint8_t* img; long long* d;
  cudaMalloc(&img, 68 * 16384); cudaMemset(img, 0, 68 * 16384); cudaMalloc(&d, 148 * 8);
  const int smem = 65536 + S * 16384 + 1024;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  for (int waiters : {0, 16})
    for (int issuers : {1, 2}) {
      const int mode = 0;
      k<<<148, 96 + 32 * waiters, smem>>>(img, d, 4000, waiters, mode, issuers);
      cudaError_t e = cudaDeviceSynchronize();
      std::vector<long long> h(148); cudaMemcpy(h.data(), d, 148 * 8, cudaMemcpyDeviceToHost);
      double s = 0; for (int i = 0; i < 148; ++i) s += h[i];
      printf("waiting warps %2d issuers %d : %.1f clk/UMMA (%s)\n", waiters, issuers, s / 148 / 100, cudaGetErrorString(e));
    }
}
