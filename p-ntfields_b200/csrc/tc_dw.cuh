// tcgen05 weight-gradient kernel (tensor-core precision mode):
//     dW_l[m][k] = sum over tiles and rows of  Ybar_l[m][row] * X_l[k][row]        (l = 0 .. 12)
// Both operands are the fp16 tile images stashed by the forward (X_l) and reverse (Ybar_l) kernels; read
// as K-major operands whose K axis is the row axis, they need no transposition (tc_common.cuh).  A CTA
// owns one work unit (layer, 128-feature half of M) over a strided subset of the tiles, accumulates the
// 128 x K_l product in TMEM over all of its tiles and adds it to the flat gradient with atomics.
// The reference gets these from autograd (loss.backward(), models/model_res_sigmoid.py:1071).
#pragma once
#include "tc_forward.cuh"

namespace pntf {
namespace tc {

constexpr int kDwMaxUnits = 19;
struct DwUnits {
  int n;
  int l[kDwMaxUnits], h[kDwMaxUnits], cta_begin[kDwMaxUnits + 1];
};

struct TcDwArgs {
  TcStash st;
  DwUnits units;
  int64_t ntiles;
  float* grad;
  float out_scale;
  const float* sched;       // optional device-resident {alpha, beta, gamma}: out_scale = sched[1] * inv_scale
  float inv_scale;
  // fused backward (tc_backward.cuh): wait for flags[tile * 16 + l] == epoch before reading Ybar_l of a tile
  const uint32_t* flags;
  uint32_t epoch;
  long long* detw;          // deterministic mode: 64-bit fixed-point accumulators [PNTF_PARAM_COUNT] (tc_reverse.cuh: det_add)
};

// acquire side of the reverse sweep's flags (tc_reverse.cuh: flag_publish)
__device__ __forceinline__ void dw_flag_wait(const uint32_t* flag, uint32_t epoch) {
  uint32_t v, spins = 0;
  for (;;) {
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(flag) : "memory");
    if (v == epoch) break;
    __nanosleep(64);
    if (++spins > (1u << 25)) __trap();   // seconds: a protocol bug traps instead of hanging the GPU
  }
  asm volatile("fence.proxy.async;" ::: "memory");
}

constexpr int kDwStageBytes = 128 * 256 + 256 * 256;   // Ybar half tile + X tile
constexpr int kDwStages = 2;
constexpr int kDwThreads = 320;                         // 8 drain warps + loader warp 8 + MMA warp 9

struct TcDwCtl {
  uint64_t full[kDwStages], empty[kDwStages], done;
  uint32_t tmem_base;
};

// `cta`: index of this CTA among the weight-gradient CTAs of the launch.  Warps 0..7 drain, warp 8 loads, warp 9 issues;
// further warps of the block (fused launch: the block has the reverse sweep's size) only take part in the barriers.
__device__ __forceinline__ void tc_dw_role(const TcDwArgs& a, const int cta) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ TcDwCtl ctl;
  const int tid = threadIdx.x, warp = tid >> 5, lane_id = tid & 31;
  int u = 0;
  for (int i = 1; i < a.units.n; ++i)
    if (cta >= a.units.cta_begin[i]) u = i;
  const int l = a.units.l[u], h = a.units.h[u];
  const int rank = cta - a.units.cta_begin[u];
  const int ncta = a.units.cta_begin[u + 1] - a.units.cta_begin[u];
  const int K = c_K[l], M = c_M[l];
  const int64_t my_tiles = (a.ntiles > rank) ? (a.ntiles - rank + ncta - 1) / ncta : 0;

  if (tid == 0) {
    for (int i = 0; i < kDwStages; ++i) { mbar_init(&ctl.full[i], 1); mbar_init(&ctl.empty[i], 1); }
    mbar_init(&ctl.done, 1);
    fence_barrier_init();
  }
  if (warp == 9) tmem_alloc(&ctl.tmem_base, 256);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = ctl.tmem_base;

  if (warp == 8) {
    if (lane_id == 0) {
      const uint32_t xbytes = uint32_t(K) * 256;
      for (int64_t i = 0; i < my_tiles; ++i) {
        const int64_t tile = rank + i * ncta;
        const uint32_t slot = uint32_t(i % kDwStages), ph = uint32_t(i / kDwStages) & 1;
        mbar_wait(&ctl.empty[slot], ph ^ 1);
        if (a.flags != nullptr) dw_flag_wait(a.flags + size_t(tile) * 16 + l, a.epoch);
        mbar_arrive_expect_tx(&ctl.full[slot], 128 * 256 + xbytes);
        uint8_t* dst = smem + slot * kDwStageBytes;
        bulk_g2s(dst, a.st.G[l] + (size_t(tile) * M + size_t(h) * 128) * 256, 128 * 256, &ctl.full[slot]);
        bulk_g2s(dst + 128 * 256, a.st.X[l] + size_t(tile) * xbytes, xbytes, &ctl.full[slot]);
      }
    }
  } else if (warp == 9) {
    if (lane_id == 0 && my_tiles > 0) {
      const uint32_t idesc = make_idesc(128, K, false, false);
      for (int64_t i = 0; i < my_tiles; ++i) {
        const uint32_t slot = uint32_t(i % kDwStages), ph = uint32_t(i / kDwStages) & 1;
        mbar_wait(&ctl.full[slot], ph);
        tc_fence_after();
        const uint32_t g_addr = smem_u32(smem + slot * kDwStageBytes), x_addr = g_addr + 128 * 256;
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {     // 128 rows = 8 steps of 16
          const uint32_t off = (ks >> 2) * kAtomBytes + (ks & 3) * 32;
          umma_f16(tmem_base, make_desc(g_addr + off, 16, kActSBO), make_desc(x_addr + off, 16, kActSBO), idesc,
                   (i > 0 || ks > 0) ? 1u : 0u);
        }
        umma_commit(&ctl.empty[slot]);
      }
      umma_commit(&ctl.done);
    }
  } else if (warp < 8 && my_tiles > 0) {
    mbar_wait(&ctl.done, 0);
    tc_fence_after();
    const int m = tid & 127, half = tid >> 7;
    const uint32_t lane_base = tmem_base + (uint32_t((warp & 3) * 32) << 16);
    float* dst = a.grad + c_w[l] + size_t(h * 128 + m) * K;
    long long* dstq = a.detw != nullptr ? a.detw + c_w[l] + size_t(h * 128 + m) * K : nullptr;
    const bool row_ok = (h * 128 + m) < M;
    const float sc = a.sched != nullptr ? __ldg(a.sched + 1) * a.inv_scale : a.out_scale;
    const int kc = K / 2;
    for (int c0 = half * kc; c0 < (half + 1) * kc; c0 += 32) {
      float v[32];
      tmem_ld<32>(lane_base + c0, v);
      if (row_ok) {
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          if (dstq != nullptr)
            atomicAdd(reinterpret_cast<unsigned long long*>(dstq + c0 + j),
                      static_cast<unsigned long long>(__float2ll_rn(v[j] * 16777216.0f)));
          else
            atomicAdd(dst + c0 + j, v[j] * sc);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 9) tmem_dealloc(tmem_base, 256);
}

// stand-alone launch (PNTF_BWD_SPLIT=1), after the reverse kernel has finished
__global__ void __launch_bounds__(kDwThreads, 1) tc_dw_kernel(const TcDwArgs a) { tc_dw_role(a, int(blockIdx.x)); }

}  // namespace tc
}  // namespace pntf
