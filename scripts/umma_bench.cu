// Micro-benchmark of the tcgen05 issue rate in the shapes the forward kernel uses (not product code).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o scripts/umma_bench.bin scripts/umma_bench.cu
//   scripts/umma_bench.bin
// Each CTA issues NU UMMAs (M=128, K=16, fp16 -> fp32) from one thread and reports clocks per UMMA:
//   mode 0: operands resident in shared memory, B MN-major (the forward kernel's activation tiles)
//   mode 1: operands resident, B K-major
//   mode 2: as 0, but the A operand is streamed from global memory (L2) through a ring of S 16 KB stages by
//           a producer thread, 4 UMMAs per stage (the forward kernel's weight ring); REUSE UMMA groups per stage
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../p-ntfields_b200/csrc/tc_common.cuh"

using namespace pntf::tc;

__device__ __forceinline__ uint32_t mbar_test(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile("{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  return ok;
}
struct Args {
  const uint8_t* wimg;
  long long* clk;
  int nu, n, b_mn, stream, ring, reuse, nstages_img, commit_every, copy_bytes, nprod, pmode, cmode;
};

__global__ void __launch_bounds__(128, 1) bench_kernel(Args a) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* act = smem;                 // 64 KB activation tile
  uint8_t* ring = smem + 65536;        // S stages
  __shared__ uint64_t full[9], empty[9], done;
  __shared__ uint32_t tmem_base_s;
  const int t = threadIdx.x, warp = t >> 5;
  if (t == 0) {
    for (int i = 0; i < 9; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    mbar_init(&done, 1);
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc(&tmem_base_s, 512);
  for (int i = t; i < (65536 + a.ring * kStageBytes) / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tb = tmem_base_s;
  const int S = a.ring;
  if ((warp == 1 || warp == 2) && a.stream) {
    const int nst = a.nu / (4 * a.reuse);
    long long pw = 0, pi = 0;
    if (a.pmode == 2) {          // two producer warps, alternate stages
      if ((t & 31) == 0) {
        for (int it = warp - 1; it < nst; it += 2) {
          const uint32_t slot = it % S, ph = (it / S) & 1;
          mbar_wait(&empty[slot], ph ^ 1);
          mbar_arrive_expect_tx(&full[slot], kStageBytes);
          bulk_g2s(ring + slot * kStageBytes, a.wimg + size_t((it + blockIdx.x * 7) % a.nstages_img) * kStageBytes, kStageBytes, &full[slot]);
        }
      }
    } else if (warp == 1 && a.pmode == 3) {   // one 32 KB copy per two stages (ring slots 2j, 2j+1 contiguous; S even)
      if ((t & 31) == 0) {
        for (int it = 0; it < nst; it += 2) {
          const uint32_t slot = it % S, ph = (it / S) & 1;
          mbar_wait(&empty[slot], ph ^ 1);
          mbar_wait(&empty[slot + 1], ph ^ 1);
          mbar_arrive_expect_tx(&full[slot], kStageBytes);
          mbar_arrive_expect_tx(&full[slot + 1], kStageBytes);
          bulk_g2s(ring + slot * kStageBytes, a.wimg + size_t((it + blockIdx.x * 7) % (a.nstages_img - 1)) * kStageBytes, kStageBytes, &full[slot]);
          bulk_g2s(ring + (slot + 1) * kStageBytes, a.wimg + size_t((it + 1 + blockIdx.x * 7) % (a.nstages_img - 1)) * kStageBytes, kStageBytes, &full[slot + 1]);
        }
      }
    } else if (warp == 1 && (t & 31) < a.nprod) {
      const int lane = t & 31;
      const uint32_t part = a.copy_bytes / a.nprod;
      for (int it = 0; it < nst; ++it) {
        const uint32_t slot = it % S, ph = (it / S) & 1;
        const long long w0 = clock64();
        if (a.pmode == 1) mbar_spin(&empty[slot], ph ^ 1); else mbar_wait(&empty[slot], ph ^ 1);
        const long long w1 = clock64();
        if (it >= S) pw += w1 - w0;
        if (lane == 0) mbar_arrive_expect_tx(&full[slot], a.copy_bytes);
        bulk_g2s(ring + slot * kStageBytes + lane * part, a.wimg + size_t((it + blockIdx.x * 7) % a.nstages_img) * kStageBytes + lane * part, part, &full[slot]);
        if (it >= S) pi += clock64() - w1;
      }
      if (lane == 0 && blockIdx.x == 0) { a.clk[202] = pw / (nst - S); a.clk[203] = pi / (nst - S); }
    }
  } else if (warp == 0) {
    const uint32_t idesc = (1u << 4) | (0u << 15) | (uint32_t(a.b_mn) << 16) | (uint32_t(a.n >> 3) << 17) | (uint32_t(128 >> 4) << 24);
    const uint32_t a_hi = desc_hi(1024), b_hi = desc_hi(a.b_mn ? kActSBO : 1024);
    const uint32_t ring_addr = smem_u32(ring), act_addr = smem_u32(act);
    long long t0 = 0, wait_c = 0, issue_c = 0;
    if (a.stream) {
      const int nst = a.nu / (4 * a.reuse);
      if (a.cmode >= 5) {
        uint32_t ok = mbar_test(&full[0], 0);
        for (int it = 0; it < nst; ++it) {
          const uint32_t slot = it % S, ph = (it / S) & 1;
          if (!ok) mbar_spin(&full[slot], ph);
          const uint32_t nslot = (it + 1) % S, nph = ((it + 1) / S) & 1;
          ok = mbar_test(&full[nslot], nph);
          if (a.cmode == 5) tc_fence_after();
          if (it == S) t0 = clock64();
          if (elect_one()) {
            const uint32_t a_lo = desc_lo(ring_addr + slot * kStageBytes, 16);
            const uint32_t b_lo = desc_lo(act_addr + (it & 3) * 8 * kActSBO, kActLBO);
#pragma unroll
            for (int j = 0; j < 4; ++j)
              umma_f16_split(tb, a_lo + 2 * j, a_hi, b_lo + j * ((2 * kActSBO) >> 4), b_hi, idesc, 1u);
            umma_commit(&empty[slot]);
          }
          __syncwarp();
        }
      } else if (a.cmode >= 2) {
        if ((t & 31) == 0) {
          for (int it = 0; it < nst; ++it) {
            const uint32_t slot = it % S, ph = (it / S) & 1;
            if (a.cmode == 4) mbar_wait(&full[slot], ph); else mbar_spin(&full[slot], ph);
            if (a.cmode != 3) tc_fence_after();
            if (it == S) t0 = clock64();
            const uint32_t a_lo = desc_lo(ring_addr + slot * kStageBytes, 16);
            const uint32_t b_lo = desc_lo(act_addr + (it & 3) * 8 * kActSBO, kActLBO);
#pragma unroll
            for (int j = 0; j < 4; ++j)
              umma_f16_split(tb, a_lo + 2 * j, a_hi, b_lo + j * ((2 * kActSBO) >> 4), b_hi, idesc, 1u);
            umma_commit(&empty[slot]);
          }
        }
        t0 = __shfl_sync(0xffffffffu, t0, 0);
        __syncwarp();
      } else
      for (int it = 0; it < nst; ++it) {
        const uint32_t slot = it % S, ph = (it / S) & 1;
        const long long w0 = clock64();
        mbar_spin(&full[slot], ph);
        const long long w1 = clock64();
        if (it >= S) wait_c += w1 - w0;
        if (a.cmode != 1) tc_fence_after();
        if (it == S) t0 = clock64();
        if (elect_one()) {
          const uint32_t a_lo = desc_lo(ring_addr + slot * kStageBytes, 16);
          for (int r = 0; r < a.reuse; ++r) {
            const uint32_t b_lo = desc_lo(act_addr + ((it + r) & 3) * 8 * kActSBO, kActLBO);
#pragma unroll
            for (int j = 0; j < 4; ++j)
              umma_f16_split(tb + (r & 1) * 256, a_lo + 2 * j, a_hi, b_lo + j * ((2 * kActSBO) >> 4), b_hi, idesc, 1u);
          }
          umma_commit(&empty[slot]);
        }
        __syncwarp();
        if (it >= S) issue_c += clock64() - w1;
      }
      if ((t & 31) == 0 && blockIdx.x == 0) { a.clk[200] = wait_c / (nst - S); a.clk[201] = issue_c / (nst - S); }
      if (elect_one()) umma_commit(&done);
      __syncwarp();
      mbar_wait(&done, 0);
      const long long t1 = clock64();
      if ((t & 31) == 0) a.clk[blockIdx.x] = (t1 - t0) * 1000 / ((nst - S) * 4 * a.reuse);
    } else {
      t0 = clock64();
      if (elect_one()) {
        for (int it = 0; it < a.nu / 4; ++it) {
          const uint32_t a_lo = desc_lo(ring_addr + (it % S) * kStageBytes, 16);
          uint32_t b_lo;
          if (a.b_mn) b_lo = desc_lo(act_addr + (it & 3) * 8 * kActSBO, kActLBO);
          else b_lo = desc_lo(act_addr + (it & 3) * 16384, 16);
#pragma unroll
          for (int j = 0; j < 4; ++j)
            umma_f16_split(tb, a_lo + 2 * j, a_hi, a.b_mn ? b_lo + j * ((2 * kActSBO) >> 4) : b_lo + 2 * j, b_hi, idesc, 1u);
          if (a.commit_every) umma_commit(&empty[it % S]);
        }
        umma_commit(&done);
      }
      __syncwarp();
      mbar_wait(&done, 0);
      const long long t1 = clock64();
      if ((t & 31) == 0) a.clk[blockIdx.x] = (t1 - t0) * 1000 / a.nu;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tb, 512);
}

int main() {
  const int nst_img = 68;
  uint8_t* wimg;
  long long* clk;
  cudaMalloc(&wimg, size_t(nst_img) * kStageBytes);
  cudaMemset(wimg, 0, size_t(nst_img) * kStageBytes);
  cudaMalloc(&clk, 256 * sizeof(long long));
  struct Cfg { int n, b_mn, stream, ring, reuse, ctas, ce, cb, np, pm, cm; };
  std::vector<Cfg> cfgs = {
      {128, 1, 1, 5, 1, 148, 0, 16384, 1, 0, 0}, {128, 1, 1, 5, 1, 148, 0, 16384, 1, 0, 5}, {128, 1, 1, 5, 1, 148, 0, 16384, 1, 0, 6},
      {128, 1, 1, 3, 1, 148, 0, 16384, 1, 0, 6}, {64, 1, 1, 5, 1, 148, 0, 16384, 1, 0, 6},
  };
  for (const Cfg& c : cfgs) {
    Args a{wimg, clk, 4096, c.n, c.b_mn, c.stream, c.ring, c.reuse, nst_img, c.ce, c.cb, c.np, c.pm, c.cm};
    const int smem = 65536 + 9 * kStageBytes + 1024;
    cudaFuncSetAttribute(bench_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    cudaMemset(clk, 0, 256 * sizeof(long long));
    bench_kernel<<<c.ctas, 128, smem>>>(a);
    cudaError_t e = cudaDeviceSynchronize();
    std::vector<long long> h(256);
    cudaMemcpy(h.data(), clk, 256 * sizeof(long long), cudaMemcpyDeviceToHost);
    double s = 0, mx = 0;
    for (int i = 0; i < c.ctas; ++i) { s += h[i]; mx = h[i] > mx ? h[i] : mx; }
    printf("N=%3d b_mn=%d stream=%d ring=%d reuse=%d ctas=%3d commit_every=%d copy=%d nprod=%d pmode=%d cmode=%d : %.1f clk/UMMA (max %.1f) %s\n", c.n, c.b_mn, c.stream, c.ring, c.reuse,
           c.ctas, c.ce, c.cb, c.np, c.pm, c.cm, s / c.ctas / 1000.0, mx / 1000.0, cudaGetErrorString(e));
    printf("    per stage: consumer wait %lld issue %lld | producer wait %lld issue %lld\n", h[200], h[201], h[202], h[203]);
  }
  return 0;
}
