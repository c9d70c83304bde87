// Self-test of the tcgen05 plumbing in tc_common.cuh (operand layouts, descriptors, TMEM mapping).
// One CTA computes, for W[128x128], X[128 features x 128 rows], C[128x128] (fp32, row-major):
//   out0 = W X            (fresh accumulator)
//   out1 = C + W X        (accumulator pre-loaded with tcgen05.st: the residual-skip path)
//   out2 = X X^T over rows, i.e. out2[m][n] = sum_r X[m][r] X[n][r]   (activation tiles used as
//          K-major operands whose K is the row axis: the dW contraction)
// with operands rounded to fp16.  tests/test_gpu_parity.py compares with numpy.
#include "../../include/pntf_b200.h"
#include "tc_common.cuh"

namespace pntf {
namespace tc {

__global__ void __launch_bounds__(128, 1) selftest_kernel(const float* __restrict__ W, const float* __restrict__ X,
                                                          const float* __restrict__ C, float* __restrict__ out) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* wst = smem;                    // 2 weight stages (K = 128)
  uint8_t* act = smem + 2 * kStageBytes;  // activation tile 128 features x 128 rows = 32 KB
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_base_s;
  const int t = threadIdx.x, warp = t >> 5;
  if (t == 0) {
    mbar_init(&bar, 1);
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc(&tmem_base_s, 512);
  // weights: thread t = out-feature m
  for (int k = 0; k < 128; ++k) {
    const __half h = __float2half_rn(W[t * 128 + k]);
    *reinterpret_cast<__half*>(wst + (k >> 6) * kStageBytes + wstage_off(t, k & 63)) = h;
  }
  // activations: thread t = feature k, 16 chunks of 8 rows
  for (int c = 0; c < 16; ++c) {
    float v[8];
    for (int i = 0; i < 8; ++i) v[i] = X[t * 128 + c * 8 + i];
    st_chunk(act, t, c >> 3, c & 7, v);
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tb = tmem_base_s;
  // pre-load accumulator 1 (columns 128..255) with C: thread t = lane m
  for (int j = 0; j < 8; ++j) {
    float v[16];
    for (int i = 0; i < 16; ++i) v[i] = C[t * 128 + j * 16 + i];
    tmem_st16(tmem_addr(tb, warp * 32, 128 + j * 16), v);
  }
  tmem_st_wait();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (t == 0) {
    const uint32_t idesc = make_idesc(128, 128, false, true);
    for (int ks = 0; ks < 8; ++ks) {   // K = 128 = 8 steps of 16
      const uint64_t ad = make_desc(smem_u32(wst) + (ks >> 2) * kStageBytes + (ks & 3) * 32, 16, 1024);
      const uint64_t bd = make_desc(smem_u32(act) + ks * 2 * kActSBO, kActLBO, kActSBO);
      umma_f16(tb + 0, ad, bd, idesc, ks > 0);
      umma_f16(tb + 128, ad, bd, idesc, 1);
    }
    // out2: both operands = activation tile, K-major over rows (8 steps of 16 rows)
    const uint32_t idesc2 = make_idesc(128, 128, false, false);
    for (int ks = 0; ks < 8; ++ks) {
      const uint64_t d = make_desc(smem_u32(act) + (ks >> 2) * kAtomBytes + (ks & 3) * 32, 16, kActSBO);
      umma_f16(tb + 256, d, d, idesc2, ks > 0);
    }
    umma_commit(&bar);
  }
  mbar_wait(&bar, 0);
  tc_fence_after();
  for (int o = 0; o < 3; ++o)
    for (int j = 0; j < 8; ++j) {
      float v[16];
      tmem_ld16(tmem_addr(tb, warp * 32, o * 128 + j * 16), v);
      for (int i = 0; i < 16; ++i) out[(o * 128 + t) * 128 + j * 16 + i] = v[i];
    }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tb, 512);
}

}  // namespace tc
}  // namespace pntf

extern "C" int pntf_tc_selftest(const float* W, const float* X, const float* C, float* out, void* stream) {
  using namespace pntf::tc;
  const int smem = 2 * kStageBytes + 32768 + 1024;
  if (cudaFuncSetAttribute(selftest_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) return PNTF_ERR_CUDA;
  selftest_kernel<<<1, 128, smem, static_cast<cudaStream_t>(stream)>>>(W, X, C, out);
  return cudaGetLastError() == cudaSuccess ? PNTF_OK : PNTF_ERR_CUDA;
}
