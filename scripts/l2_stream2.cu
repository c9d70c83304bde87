// Where do the ~400 clk per cp.async.bulk of a single producer thread go?  Variants of the producer loop.
#include <cstdio>
#include <vector>
#include "../p-ntfields_b200/csrc/tc_common.cuh"
using namespace pntf::tc;
__global__ void __launch_bounds__(64, 1) k(const uint8_t* img, long long* out, int nst, int mode) {
  extern __shared__ __align__(1024) uint8_t smem[];
  constexpr int S = 6, bytes = 16384;
  __shared__ uint64_t full[S], empty[S];
  const int t = threadIdx.x;
  if (t == 0) { for (int i = 0; i < S; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); } fence_barrier_init(); }
  __syncthreads();
  if (t == 0) {
    if (mode == 0) {
      for (int it = 0; it < nst; ++it) {
        const uint32_t slot = it % S, ph = (it / S) & 1;
        mbar_wait(&empty[slot], ph ^ 1);
        mbar_arrive_expect_tx(&full[slot], bytes);
        bulk_g2s(smem + slot * bytes, img + size_t(it % 68) * 16384, bytes, &full[slot]);
      }
    } else if (mode == 1) {   // one barrier round, two 8 KB copies
      for (int it = 0; it < nst; ++it) {
        const uint32_t slot = it % S, ph = (it / S) & 1;
        mbar_wait(&empty[slot], ph ^ 1);
        mbar_arrive_expect_tx(&full[slot], bytes);
        bulk_g2s(smem + slot * bytes, img + size_t(it % 68) * 16384, bytes / 2, &full[slot]);
        bulk_g2s(smem + slot * bytes + bytes / 2, img + size_t(it % 68) * 16384 + bytes / 2, bytes / 2, &full[slot]);
      }
    } else if (mode == 2) {   // unrolled by two, grouped: waits, expects, copies
      for (int it = 0; it < nst; it += 2) {
        const uint32_t s0 = it % S, p0 = (it / S) & 1, s1 = (it + 1) % S, p1 = ((it + 1) / S) & 1;
        mbar_wait(&empty[s0], p0 ^ 1);
        mbar_wait(&empty[s1], p1 ^ 1);
        mbar_arrive_expect_tx(&full[s0], bytes);
        mbar_arrive_expect_tx(&full[s1], bytes);
        bulk_g2s(smem + s0 * bytes, img + size_t(it % 68) * 16384, bytes, &full[s0]);
        bulk_g2s(smem + s1 * bytes, img + size_t((it + 1) % 68) * 16384, bytes, &full[s1]);
      }
    } else if (mode == 3) {   // spin (test_wait) instead of try_wait
      for (int it = 0; it < nst; ++it) {
        const uint32_t slot = it % S, ph = (it / S) & 1;
        mbar_spin(&empty[slot], ph ^ 1);
        mbar_arrive_expect_tx(&full[slot], bytes);
        bulk_g2s(smem + slot * bytes, img + size_t(it % 68) * 16384, bytes, &full[slot]);
      }
    } else if (mode == 4) {   // unrolled by three
      for (int it = 0; it < nst; it += 3) {
        uint32_t s[3], p[3];
        for (int j = 0; j < 3; ++j) { s[j] = (it + j) % S; p[j] = ((it + j) / S) & 1; }
        for (int j = 0; j < 3; ++j) mbar_wait(&empty[s[j]], p[j] ^ 1);
        for (int j = 0; j < 3; ++j) mbar_arrive_expect_tx(&full[s[j]], bytes);
        for (int j = 0; j < 3; ++j) bulk_g2s(smem + s[j] * bytes, img + size_t((it + j) % 68) * 16384, bytes, &full[s[j]]);
      }
    }
  } else if (t == 32) {
    long long t0 = 0;
    for (int it = 0; it < nst; ++it) {
      const uint32_t slot = it % S, ph = (it / S) & 1;
      mbar_wait(&full[slot], ph);
      if (it == S) t0 = clock64();
      mbar_arrive(&empty[slot]);
    }
    out[blockIdx.x] = clock64() - t0;
  }
}
int main() {
  uint8_t* img; long long* d;
  cudaMalloc(&img, 69 * 16384); cudaMemset(img, 1, 69 * 16384); cudaMalloc(&d, 148 * 8);
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 6 * 16384);
  const int nst = 6000;
  for (int mode : {0, 1, 2, 3, 4}) {
    k<<<148, 64, 6 * 16384>>>(img, d, nst, mode);
    cudaError_t e = cudaDeviceSynchronize();
    std::vector<long long> h(148); cudaMemcpy(h.data(), d, 148 * 8, cudaMemcpyDeviceToHost);
    double mx = 0; for (int i = 0; i < 148; ++i) mx = h[i] > mx ? h[i] : mx;
    printf("mode %d : %.1f B/clk/SM, %.0f clk per 16 KB (%s)\n", mode, double(nst - 6) * 16384 / mx, mx / (nst - 6), cudaGetErrorString(e));
  }
}
