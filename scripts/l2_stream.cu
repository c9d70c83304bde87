// L2 -> shared-memory streaming ceiling with cp.async.bulk (the weight ring's traffic pattern): every CTA streams
// the same 1.06 MB image (68 x 16 KB stages) round and round through a ring of S slots; a consumer thread releases
// each slot as soon as it lands.  Reports bytes / clk / SM and the chip-wide TB/s.
#include <cstdio>
#include <vector>
#include "../p-ntfields_b200/csrc/tc_common.cuh"
using namespace pntf::tc;
__global__ void __launch_bounds__(256, 1) k(const uint8_t* img, long long* out, int nst, int S, int bytes, int np) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t fullb[4][12], emptyb[4][12];
  const int t = threadIdx.x;
  if (t == 0) { for (int q = 0; q < 4; ++q) for (int i = 0; i < 12; ++i) { mbar_init(&fullb[q][i], 1); mbar_init(&emptyb[q][i], 1); } fence_barrier_init(); }
  __syncthreads();
  const int q = (t >> 6) & 3;   // producer/consumer pair q uses warps 2q, 2q+1
  uint64_t* full = fullb[q]; uint64_t* empty = emptyb[q];
  if (q < np && (t & 63) == 0) {
    for (int it = 0; it < nst; ++it) {
      const uint32_t slot = it % S, ph = (it / S) & 1;
      mbar_wait(&empty[slot], ph ^ 1);
      mbar_arrive_expect_tx(&full[slot], bytes);
      bulk_g2s(smem + (q * S + slot) * bytes, img + size_t((it + blockIdx.x * 5 + q * 17) % 68) * 16384, bytes, &full[slot]);
    }
  } else if (q < np && (t & 63) == 32) {
    long long t0 = 0;
    for (int it = 0; it < nst; ++it) {
      const uint32_t slot = it % S, ph = (it / S) & 1;
      mbar_wait(&full[slot], ph);
      if (it == S) t0 = clock64();
      mbar_arrive(&empty[slot]);
    }
    if (q == 0) out[blockIdx.x] = clock64() - t0;
  }
}
int main() {
  uint8_t* img; long long* d;
  cudaMalloc(&img, 68 * 16384 + 65536); cudaMemset(img, 1, 68 * 16384 + 65536); cudaMalloc(&d, 148 * 8);
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  const int nst = 4000;
  for (int ctas : {1, 148})
    for (int np : {1, 2, 4})
      for (int bytes : {16384, 8192}) {
        const int S = 3;
        k<<<ctas, 256, 4 * S * bytes>>>(img, d, nst, S, bytes, np);
        cudaError_t e = cudaDeviceSynchronize();
        std::vector<long long> h(148); cudaMemcpy(h.data(), d, ctas * 8, cudaMemcpyDeviceToHost);
        double mx = 0; for (int i = 0; i < ctas; ++i) mx = h[i] > mx ? h[i] : mx;
        const double bpc = double(nst - S) * bytes * np / mx;
        printf("ctas=%3d producers=%d ring=%2d x %d B : %.1f B/clk/SM, %.1f KB/clk chip (%s)\n", ctas, np, S, bytes, bpc, bpc * ctas / 1024, cudaGetErrorString(e));
      }
}
