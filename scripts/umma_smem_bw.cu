// Does shared-memory traffic from OTHER agents slow the tensor pipe?  One warp issues N=128 (or N=64 / N=256) UMMAs back to
// back from resident operands (no barriers inside the loop); optionally P producer threads stream 16 KB bulk copies
// L2 -> shared memory into a separate ring that nobody consumes, and W warps stream st.shared into a third region.
// Reports clk per UMMA and the producers' achieved B/clk.  Model under test: operands + fill + stores share 128 B/clk.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o scripts/umma_smem_bw.bin scripts/umma_smem_bw.cu
#include <cstdio>
#include <vector>
#include "../p-ntfields_b200/csrc/tc_common.cuh"
using namespace pntf::tc;
constexpr int S = 4;             // copies in flight per producer thread
constexpr int kMaxProd = 4;
__global__ void __launch_bounds__(768, 1) k(const uint8_t* img, long long* out, int numma, int n_rows, int producers, int writers, int commit_every, int tmem_readers = 0) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* act = smem;                       // 64 KB operand tile (B)
  uint8_t* wts = smem + 65536;               // 16 KB weight stage (A)
  uint8_t* ring = smem + 65536 + 16384;      // kMaxProd * S * 16 KB fill target... limited to 4 slots total below
  uint8_t* scratch = ring + 4 * 16384;       // 16 KB st.shared target
  __shared__ uint64_t full[kMaxProd][S], done, dummy;
  __shared__ uint32_t tmem_base_s;
  __shared__ volatile int stop;
  const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
  if (t == 0) {
    for (int p = 0; p < kMaxProd; ++p) for (int i = 0; i < S; ++i) mbar_init(&full[p][i], 1);
    mbar_init(&done, 1); mbar_init(&dummy, 1); stop = 0; fence_barrier_init();
  }
  if (warp == 0) tmem_alloc(&tmem_base_s, 512);
  for (int i = t; i < (65536 + 16384) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  fence_proxy_async(); tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tb = tmem_base_s;
  if (warp == 0) {
    const uint32_t idesc = make_idesc(128, n_rows, false, true);
    constexpr uint32_t a_hi = desc_hi(1024), b_hi = desc_hi(kActSBO);
    const uint32_t a_lo = desc_lo(smem_u32(wts), 16), b_lo0 = desc_lo(smem_u32(act), kActLBO);
    long long t0 = clock64();
    if (elect_one()) {
      for (int it = 0; it < numma; it += 4) {
        const uint32_t b_lo = b_lo0 + ((it >> 2) & 3) * ((8 * kActSBO) >> 4);
#pragma unroll
        for (int j = 0; j < 4; ++j) umma_f16_split(tb, a_lo + 2 * j, a_hi, b_lo + j * ((2 * kActSBO) >> 4), b_hi, idesc, 1u);
        if (commit_every && ((it + 4) % commit_every) == 0) umma_commit(&dummy);   // a commit nobody waits for
      }
      umma_commit(&done);
    }
    __syncwarp();
    mbar_wait(&done, 0);
    const long long t1 = clock64();
    if (lane == 0) { out[blockIdx.x * 2] = (t1 - t0) * 100 / numma; stop = 1; }
  } else if (warp == 1) {
    // producers: lane p streams 16 KB copies into slot p of the ring (S in flight on S barriers), never consumed
    if (lane < producers) {
      long long n = 0; const long long t0 = clock64();
      uint32_t ph[S] = {0, 0, 0, 0};
      for (int i = 0; i < S; ++i) {
        mbar_arrive_expect_tx(&full[lane][i], kStageBytes);
        bulk_g2s(ring + lane * kStageBytes, img + size_t((n++) % 68) * kStageBytes, kStageBytes, &full[lane][i]);
      }
      while (!stop) {
        for (int i = 0; i < S; ++i) {
          mbar_wait(&full[lane][i], ph[i]); ph[i] ^= 1;
          mbar_arrive_expect_tx(&full[lane][i], kStageBytes);
          bulk_g2s(ring + lane * kStageBytes, img + size_t((n++) % 68) * kStageBytes, kStageBytes, &full[lane][i]);
        }
      }
      for (int i = 0; i < S; ++i) mbar_wait(&full[lane][i], ph[i]);
      const long long t1 = clock64();
      if (lane == 0) out[blockIdx.x * 2 + 1] = n * producers * kStageBytes * 100 / (t1 - t0);
    }
  } else if (warp >= 2 && warp < 2 + tmem_readers) {
    // epilogue-like TMEM traffic: tcgen05.ld of 32 columns + tcgen05.st of 32 columns per iteration, in the half of
    // TMEM the UMMAs do not touch (columns 256..511)
    float v[32];
    const uint32_t taddr = tb + (uint32_t((warp & 3) * 32) << 16) + 256 + 32 * ((warp >> 2) & 3);
    float acc = 0.f;
    while (!stop) {
      tmem_ld<32>(taddr, v);
#pragma unroll
      for (int i = 0; i < 32; ++i) { acc += v[i]; v[i] = acc; }
      tmem_st<32>(taddr, v);
      tmem_st_wait();
    }
    if (acc == 123.456f) out[0] = 1;
  } else if (warp >= 2 && warp < 2 + writers) {
    // st.shared streams: 16 B per thread per store, conflict-free
    uint4 v = make_uint4(t, t, t, t);
    uint4* dst = reinterpret_cast<uint4*>(scratch) + (t & 1023);
    while (!stop) {
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(smem_u32(dst)), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
        v.x += i;
      }
    }
  }
  tc_fence_before(); __syncthreads();
  if (warp == 0) tmem_dealloc(tb, 512);
}
int main() {
  uint8_t* img; long long* d;
  cudaMalloc(&img, 68 * 16384); cudaMemset(img, 0, 68 * 16384); cudaMalloc(&d, 148 * 16);
  const int smem = 65536 + 16384 + 4 * 16384 + 16384 + 1024;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  for (int tmem_readers : {0, 4, 8, 16}) {
    cudaMemset(d, 0, 148 * 16);
    k<<<148, 64 + 32 * 16, smem>>>(img, d, 8000, 128, 1, 0, 4, tmem_readers);
    cudaError_t e = cudaDeviceSynchronize();
    std::vector<long long> h(296); cudaMemcpy(h.data(), d, 148 * 16, cudaMemcpyDeviceToHost);
    double s = 0; for (int i = 0; i < 148; ++i) s += h[2 * i];
    printf("N=128, commit every 4, 1 producer, %2d warps of tcgen05.ld+st (32 columns) beside it : %6.1f clk/UMMA (%s)\n",
           tmem_readers, s / 148 / 100, cudaGetErrorString(e));
  }
  for (int n_rows : {128})
    for (int commit_every : {0, 4, 8, 16})
      for (int producers : {0, 1}) {
        const int writers = 0;
        cudaMemset(d, 0, 148 * 16);
        k<<<148, 64 + 32 * 16, smem>>>(img, d, 8000, n_rows, producers, writers, commit_every);
        cudaError_t e = cudaDeviceSynchronize();
        std::vector<long long> h(296); cudaMemcpy(h.data(), d, 148 * 16, cudaMemcpyDeviceToHost);
        double s = 0, b = 0; for (int i = 0; i < 148; ++i) { s += h[2 * i]; b += h[2 * i + 1]; }
        printf("N=%3d commit every %2d UMMAs, producers %d : %6.1f clk/UMMA, fill %5.1f B/clk (%s)\n", n_rows, commit_every, producers,
               s / 148 / 100, b / 148 / 100, cudaGetErrorString(e));
      }
  for (int n_rows : {64, 128, 256})
    for (int producers : {0, 1, 2, 4})
      for (int writers : {0, 8}) {
        cudaMemset(d, 0, 148 * 16);
        k<<<148, 64 + 32 * 16, smem>>>(img, d, 8000, n_rows, producers, writers, 0);
        cudaError_t e = cudaDeviceSynchronize();
        std::vector<long long> h(296); cudaMemcpy(h.data(), d, 148 * 16, cudaMemcpyDeviceToHost);
        double s = 0, b = 0; for (int i = 0; i < 148; ++i) { s += h[2 * i]; b += h[2 * i + 1]; }
        const double clk = s / 148 / 100, fill = b / 148 / 100;
        const double opnd = (4096.0 + n_rows * 32.0) / clk;
        printf("N=%3d producers %d st.shared warps %d : %6.1f clk/UMMA  operands %5.1f B/clk + fill %5.1f B/clk = %5.1f  (%s)\n",
               n_rows, producers, writers, clk, opnd, fill, opnd + fill, cudaGetErrorString(e));
      }
}
