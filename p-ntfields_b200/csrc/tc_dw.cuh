// tcgen05 weight-gradient kernel (tensor-core precision mode):
//     dW_l[m][k] = sum over tiles and rows of  Ybar_l[m][row] * X_l[k][row]        (l = 0 .. 12)
// Both operands are the fp16 tile images stashed by the forward (X_l) and reverse (Ybar_l) kernels; read
// as K-major operands whose K axis is the row axis, they need no transposition (tc_common.cuh).  A CTA
// owns one work unit (layer, 128-feature half of M) over a strided subset of the tiles, accumulates the
// 128 x K_l product in TMEM over all of its tiles and adds it to the flat gradient with atomics.
// Build switch -DPNTF_STASH_REGEN (tc_layers.cuh: kStashRegen, off by default -- measured slower): two operand tiles are
// then not in the stash; the CTAs of layer 0 regenerate the Fourier tile X_0 from the tile's coordinates and those of
// layer 6 the combine output X_6 from the stashed encoder outputs Y5, sixteen warps (the drain warps and eight more,
// idle until the end of the launch otherwise) writing the operand image straight into the stage.
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

// sources of the regenerated operand tiles (the training forward's arguments)
struct TcRegen {
  const float* B;           // [E][D][128]
  const float* coords;
  const int32_t* index;     // optional batch gather (pair_math.cuh: src_row)
  int cstride;
  int64_t n_total, n_per_env, pair_begin, pair_count;
};

struct TcDwArgs {
  TcStash st;
  TcRegen rg;
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
// 8 drain warps + loader warp 8 + MMA warp 9 (+ 8 generator warps 10..17 with kStashRegen)
constexpr int kDwThreads = kStashRegen ? 576 : 320;

// Operand generators: the drain warps 0..7 and the warps 10..17 (idle in the CTAs of the other layers).  Eight warps were
// not enough: two warps per scheduler of dependent transcendental chains took 3 300 clk per tile and made the units of
// layers 0 and 6 the stragglers of the launch (27.7 instead of 17.0 ms per step).
constexpr int kDwGenWarps = 16, kDwGenThreads = kDwGenWarps * 32;

struct TcDwCtl {
  uint64_t full[kDwStages], empty[kDwStages], done;
  uint64_t yfull[kDwStages];   // layer 6: the stashed encoder outputs Y5 of the slot's tile have landed
  uint32_t tmem_base;
};

// X_6 rows of feature f (smooth-max) and 128 + f (smooth-min), row quarter rq, from the Y5 image that the loader put
// into the first half of the X area (features 0..127); a thread overwrites only the lines it has read.  The forward
// combines the fp32 encoder outputs, this takes their fp16 stash image: the operand differs by fp16 rounding of the
// inputs (2^-11 relative), the same order as the operand rounding itself.
template <class C>
__device__ __forceinline__ void regen_combine_rows(uint32_t x_addr, int f, int rq) {
  constexpr int ST = C::STRIDE, EO = C::EOFF, UPT = C::UPT;
  EpiPos p;
  p.lane = f; p.rq = rq; p.nb = rq >> 1; p.cb = (rq & 1) * 4; p.tmem = 0;
  float v[32], o[32];
  ld_rows32(x_addr, p, f, v);
#pragma unroll
  for (int c = 0; c < 32; ++c) o[c] = 0.f;
#pragma unroll
  for (int u = 0; u < UPT; ++u) combine_unit<C, true>(v + u * ST, v + u * ST + EO, o + u * ST);
  st_rows32(x_addr, p, f, o);
#pragma unroll
  for (int u = 0; u < UPT; ++u) combine_unit<C, false>(v + u * ST, v + u * ST + EO, o + u * ST);
  st_rows32(x_addr, p, 128 + f, o);
}

// `cta`: index of this CTA among the weight-gradient CTAs of the launch.  Warps 0..7 drain, warp 8 loads, warp 9 issues,
// warps 0..7 and 10..17 regenerate operands in the CTAs of layers 0 and 6; further warps of the block (fused launch: the
// block has the reverse sweep's size) only take part in the barriers.
template <int D, int ORDER>
__device__ __forceinline__ void tc_dw_role(const TcDwArgs& a, const int cta) {
  using C = TcCfg<D, ORDER>;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ TcDwCtl ctl;
  __shared__ TcTile<C> gts[2];   // layer 0: coordinates of the tiles being regenerated (alternating)
  const int tid = threadIdx.x, warp = tid >> 5, lane_id = tid & 31;
  int u = 0;
  for (int i = 1; i < a.units.n; ++i)
    if (cta >= a.units.cta_begin[i]) u = i;
  const int l = a.units.l[u], h = a.units.h[u];
  const int rank = cta - a.units.cta_begin[u];
  const int ncta = a.units.cta_begin[u + 1] - a.units.cta_begin[u];
  const int K = c_K[l], M = c_M[l];
  const int64_t my_tiles = (a.ntiles > rank) ? (a.ntiles - rank + ncta - 1) / ncta : 0;
  const bool gen0 = stash_regen_layer(l) && l == 0, gen6 = stash_regen_layer(l) && l == 6, gen = gen0 || gen6;

  if (tid == 0) {
    for (int i = 0; i < kDwStages; ++i) {
      mbar_init(&ctl.full[i], gen ? 1 + kDwGenWarps : 1);   // loader (+ bytes) and, for a regenerated operand, one arrival per warp
      mbar_init(&ctl.empty[i], 1);
      mbar_init(&ctl.yfull[i], 1);
    }
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
      // A CTA keeps two stages in flight, so a tile costs it at least half a DRAM round trip under load (~2 900 clk),
      // however few bytes it loads.  The regenerating units load a third of the bytes of the others and must not take
      // as many SMs: their loader pulls the inputs of the tile kDwPrefetch positions ahead into L2, which leaves them
      // with the generator's time per tile instead.
      constexpr int64_t kDwPrefetch = 4;
      auto prefetch = [&](int64_t i) {
        if (!gen || i >= my_tiles) return;
        const int64_t tile = rank + i * ncta;
        bulk_prefetch_l2(a.st.G[l] + (size_t(tile) * M + size_t(h) * 128) * 256, 128 * 256);
        if (gen6) bulk_prefetch_l2(a.st.Y5 + size_t(tile) * kHalfTileBytes, kHalfTileBytes);
      };
      for (int64_t i = kDwStages; i < kDwPrefetch; ++i) prefetch(i);
      for (int64_t i = 0; i < my_tiles; ++i) {
        prefetch(i + kDwPrefetch);
        const int64_t tile = rank + i * ncta;
        const uint32_t slot = uint32_t(i % kDwStages), ph = uint32_t(i / kDwStages) & 1;
        mbar_wait(&ctl.empty[slot], ph ^ 1);
        if (a.flags != nullptr) dw_flag_wait(a.flags + size_t(tile) * 16 + l, a.epoch);
        mbar_arrive_expect_tx(&ctl.full[slot], 128 * 256 + (gen ? 0u : xbytes));
        uint8_t* dst = smem + slot * kDwStageBytes;
        bulk_g2s(dst, a.st.G[l] + (size_t(tile) * M + size_t(h) * 128) * 256, 128 * 256, &ctl.full[slot]);
        if (!gen) bulk_g2s(dst + 128 * 256, a.st.X[l] + size_t(tile) * xbytes, xbytes, &ctl.full[slot]);
        if (gen6) {
          mbar_arrive_expect_tx(&ctl.yfull[slot], kHalfTileBytes);
          bulk_g2s(dst + 128 * 256, a.st.Y5 + size_t(tile) * kHalfTileBytes, kHalfTileBytes, &ctl.yfull[slot]);
        }
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
  } else if ((warp < 8 || (warp >= 10 && warp < 18)) && my_tiles > 0) {
    if (gen) {
      // ----------------------------------------------------------- operand generators: thread = (feature / frequency, row quarter)
      const int gt = warp < 8 ? tid : tid - 64;     // 0..511
      const int f = gt & 127, rq = gt >> 7;
      const TcRegen& g = a.rg;
      constexpr int P = C::P;
      // coordinates of the next tile travel in registers of the first P threads (one tile ahead of their use)
      float nx[2 * D];
      int nvalid = 0, nenv = 0;
      auto fetch = [&](int64_t tile) {
        const int64_t pair = g.pair_begin + tile * P + tid;
        nvalid = (tile < a.ntiles && pair < g.pair_begin + g.pair_count && pair < g.n_total) ? 1 : 0;
        nenv = nvalid ? int(pair / g.n_per_env) : 0;
        const int64_t row = nvalid ? src_row(g, pair) : 0;
#pragma unroll
        for (int i = 0; i < 2 * D; ++i) nx[i] = nvalid ? __ldg(g.coords + row * g.cstride + i) : 0.f;
      };
      if (gen0 && tid < P) fetch(rank);
      for (int64_t i = 0; i < my_tiles; ++i) {
        const uint32_t slot = uint32_t(i % kDwStages), ph = uint32_t(i / kDwStages) & 1;
        mbar_wait(&ctl.empty[slot], ph ^ 1);   // the products of the slot's previous tile are done reading it
        const uint32_t x_addr = smem_u32(smem + slot * kDwStageBytes) + 128 * 256;
        if (gen0) {
          TcTile<C>& t = gts[i & 1];
          if (tid < P) {
#pragma unroll
            for (int k = 0; k < 2 * D; ++k) t.x[tid][k] = nx[k];
            t.valid[tid] = nvalid;
            t.env[tid] = nenv;
            fetch(rank + (i + 1) * ncta);
          }
          named_bar_sync(2, kDwGenThreads);     // gts[i & 1] is next written two tiles later, behind the barrier of tile i + 1
#ifndef PNTF_EXP_DW_NOGEN
          fourier_rows<C>(g.B, t, f, rq, x_addr);
#endif
        } else {
          mbar_wait(&ctl.yfull[slot], ph);
#ifndef PNTF_EXP_DW_NOGEN
          regen_combine_rows<C>(x_addr, f, rq);
#endif
        }
        fence_proxy_async();                    // generic-proxy writes -> visible to the UMMA operand reads
        __syncwarp();
        if (lane_id == 0) mbar_arrive(&ctl.full[slot]);
      }
    }
    if (warp < 8) {                                  // (the generator-only warps 10..17 do not drain)
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
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 9) tmem_dealloc(tmem_base, 256);
}

// stand-alone launch (PNTF_BWD_SPLIT=1), after the reverse kernel has finished
template <int D, int ORDER>
__global__ void __launch_bounds__(kDwThreads, 1) tc_dw_kernel(const TcDwArgs a) { tc_dw_role<D, ORDER>(a, int(blockIdx.x)); }

}  // namespace tc
}  // namespace pntf
