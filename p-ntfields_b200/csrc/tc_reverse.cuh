// tcgen05 reverse kernel (tensor-core precision mode): adjoint sweep of the second-order forward pass.
// Same CTA organisation as tc_forward_kernel (16 epilogue warps, producer warp, MMA warp; operand tiles
// and accumulators handed over per 128-feature half).  Per tile it seeds the adjoints from the Eikonal
// loss, then walks the layers backwards: the epilogue warps read Xbar_{l+1} (adjoint of layer l's
// activated output) from TMEM and the stashed forward outputs X_{l+1} = act(Y_l) from shared memory,
// apply the activation adjoint (with the gates recovered from the activated
// values: s = 1 - exp(-10 a), dy = dx'/s, ...), write Ybar_l as the fp16 operand tile of the next product
// Xbar_l = W_l^T Ybar_l (tcgen05.mma, transposed weight stages) and pre-load it in fp32 into the TMEM
// accumulator that receives the residual-skip adjoint.  Every Ybar_l tile is also dumped to the stash
// for the weight-gradient contraction (tc_dw_kernel); bias gradients and the 128 -> 1 output layer's
// weight gradient are reduced in shared memory.
// TMEM regions: products of the second layer of a block (and layer 12) land in B / A alternately so that
// a product never overwrites what the running epilogue still reads:
//   12 -> B ; 11 -> A ; 10 -> B (+skip) ; 9 -> A ; 8 -> B (+skip) ; 7 -> A ; 6 -> B (+skip) ; 5 -> A ;
//   4 -> A ; 3 -> B (+skip) ; 2 -> A ; 1 -> B (+skip)
// The reference gets all of this from autograd (loss.backward(), models/model_res_sigmoid.py:1071).
#pragma once
#include "tc_forward.cuh"

namespace pntf {
namespace tc {

constexpr float kGradScale = 16.0f;   // adjoints travel as fp16 scaled by this factor

struct TcRevArgs {
  TcFwdArgs f;
  const uint8_t* rimg;      // reverse (transposed) stage images
  float* grad;              // flat parameter gradient (accumulated with atomics)
  float out_scale;          // beta / (E*N) / kGradScale
  float inv_scale;          // 1 / (E*N) / kGradScale: with a device-resident schedule out_scale = sched[1] * inv_scale
  // Fused backward (tc_backward_kernel): flags[tile * kFlagStride + l] = epoch once Ybar_l of that tile is complete in
  // global memory; the weight-gradient CTAs of the same launch wait on it.  nullptr: separate launches, no flags.
  uint32_t* flags;
  uint32_t epoch;
  // Deterministic mode (pntf_set_deterministic): bias / output-layer gradient partial sums go to this CTA's row of
  // detb[cta][kDetSlots] as 64-bit fixed point (integer adds commute: the result does not depend on the order in which
  // warps and CTAs arrive), instead of fp32 adds in shared memory + fp32 atomics; tc_path.cu: det_finalize_kernel.
  long long* detb;
};

constexpr int kDetSlots = 2432 + 129;          // kNumBiasSlots + dW13[128] + db13
constexpr float kDetScale = 16777216.0f;        // 2^24 fractional bits
__device__ __forceinline__ void det_add(long long* slot, float v) {
  atomicAdd(reinterpret_cast<unsigned long long*>(slot), static_cast<unsigned long long>(__float2ll_rn(v * kDetScale)));
}

constexpr int kFlagStride = 16;   // flag words per tile (13 layers used)

// cross-CTA hand-over of bulk-stored data: the writer has waited for its bulk groups (cp.async.bulk.wait_group), the
// reader issues bulk loads after the acquire; the proxy fences order the async-proxy accesses against the flag
__device__ __forceinline__ void flag_publish(uint32_t* flag, uint32_t epoch) {
  asm volatile("fence.proxy.async;" ::: "memory");
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(flag), "r"(epoch) : "memory");
}

// accumulator region of product l (1 = B) and the region the epilogue of step l reads
static __constant__ int c_rev_acc[kNumLayers] = {0, 1, 0, 1, 0, 0, 1, 0, 1, 0, 1, 0, 1, 0};

// Adjoint of act_unit, in place on g (incoming: adjoints of the activated outputs x = (a, dx', lx') as
// stashed by the forward pass).  With t = exp(-10 a) = 1 - s:  s'/s = 10 t,  s''/s' = 10 (1 - 2 s),
// dy = dx'/s,  dy^2 s' = dy dx' 10 t,  ly s' = 10 t (lx' - dy^2 s').
template <int D, bool GEN, int ORDER>
__device__ __forceinline__ void act_adjoint_unit(float* g, const float* x) {
  constexpr int NDER = GEN ? 2 * D : D;
  const float t = ex2_approx(-14.4269504f * x[0]);      // exp(-10 a)
  const float s = 1.0f - t;
  const float r = __fdividef(1.0f, fmaxf(s, 1e-6f));
  const float t10 = 10.0f * t;
  const float c2 = fmaf(20.0f, t, -10.0f);               // 10 (1 - 2 s)
  float xb = g[0] * s;
  if constexpr (ORDER == kOrderLapSum) {
    // summed second-order column per endpoint: L' = s' Q + s L with Q = sum_i dy_i^2 (forward: act_unit)
#pragma unroll
    // with dy = dx'/s and s' = 10 s t:  dy s' = t10 dx',  Q s' = (t10 / s) sum dx'^2;  the per-coordinate work is two
    // accumulations and the derivative adjoint, the factors that do not depend on the coordinate are applied once
    const float rt10 = r * t10 * kLapScale;   // forward: Lt' = kLapScale s' Q + s Lt
    for (int e = 0; e < (GEN ? 2 : 1); ++e) {
      const float gl = g[1 + NDER + e], lxp = x[1 + NDER + e];
      const float glk = 2.0f * kLapScale * t10 * gl;
      float q = 0.f, v = 0.f;
#pragma unroll
      for (int i = 0; i < D; ++i) {
        const float dxp = x[1 + e * D + i], gd = g[1 + e * D + i];
        q = fmaf(dxp, dxp, q);
        v = fmaf(gd, dxp, v);
        g[1 + e * D + i] = fmaf(dxp, glk, gd * s);
      }
      const float u = rt10 * q;                 // kLapScale Q s'
      xb = fmaf(v, t10, xb);
      xb = fmaf(gl, fmaf(u, c2, t10 * (lxp - u)), xb);
      g[1 + NDER + e] = gl * s;
    }
  } else {
#pragma unroll
    for (int i = 0; i < NDER; ++i) {
      const float dxp = x[1 + i], lxp = x[1 + NDER + i];
      const float gd = g[1 + i], gl = g[1 + NDER + i];
      const float d10 = dxp * t10;            // dy s'
      const float u = kLapScale * dxp * r * d10;   // kLapScale dy^2 s'
      xb = fmaf(gd, d10, xb);
      xb = fmaf(gl, fmaf(u, c2, t10 * (lxp - u)), xb);
      g[1 + i] = fmaf(gd, s, 2.0f * kLapScale * gl * d10);
      g[1 + NDER + i] = gl * s;
    }
  }
  g[0] = xb;
}


// Ybar_l for one 128-feature half of a hidden layer, the thread's 32 rows.  SRC13: the incoming adjoint
// is the rank-1 product w13[f] * g13[row] (output layer) instead of a TMEM accumulator; its weight
// gradient is reduced here as well.  Returns the thread's bias-gradient partial sum.
template <class C, bool ENC, bool PRELOAD, bool SRC13>
__device__ __forceinline__ float epi_adjoint_half(const EpiPos& p, int in_col, int pre_col, int f, uint32_t x_addr,
                                                  uint32_t gt_addr, const float* g13s, float w13f, float& dw13) {
  constexpr int ST = C::STRIDE, EO = C::EOFF;
  float g[32], x[32];
  if constexpr (SRC13) {
#pragma unroll
    for (int c = 0; c < 32; ++c) g[c] = w13f * g13s[32 * p.rq + c];
  } else {
    tmem_ld<32>(p.tmem + in_col + 32 * p.rq, g);
  }
  ld_rows32(x_addr, p, f, x);
  if constexpr (SRC13) {
    // X13 = act(Y12) is the stashed tile itself: dW13[f] += sum_rows g13[row] X13[f][row]
#pragma unroll
    for (int c = 0; c < 32; ++c) dw13 = fmaf(g13s[32 * p.rq + c], x[c], dw13);
  }
  float db = 0.f;
#pragma unroll
  for (int u = 0; u < C::UPT; ++u) {
    if constexpr (ENC) {
      act_adjoint_unit<C::D, false, C::ORDER>(g + u * ST, x + u * ST);
      act_adjoint_unit<C::D, false, C::ORDER>(g + u * ST + EO, x + u * ST + EO);
      db += g[u * ST] + g[u * ST + EO];
    } else {
      act_adjoint_unit<C::D, true, C::ORDER>(g + u * ST, x + u * ST);
      db += g[u * ST];
    }
  }
  st_rows32(gt_addr, p, f, g);
  if constexpr (PRELOAD) {
    tmem_st<32>(p.tmem + pre_col + 32 * p.rq, g);
    tmem_st_wait();
  }
  return db;
}

// Adjoint of the symmetric combine of one pair (S:724-787): gmax / gmin = adjoints
// of the smooth-max / smooth-min outputs (CGP columns each), y = stashed encoder outputs Y5 of both
// endpoints (CEP columns each); writes the adjoints of the two encoder outputs to o (CGP columns).
template <class C>
__device__ __forceinline__ void combine_adjoint_unit(const float* gmax, const float* gmin, const float* y, float* o) {
  constexpr int D = C::D;
#pragma unroll
  for (int c = 0; c < C::STRIDE; ++c) o[c] = 0.f;
  const float* u = y;
  const float* v = y + C::EOFF;
  float* ou = o;
  float* ov = o + C::EOFF;
  const float s0 = fast_sigmoid(10.0f * (u[0] - v[0]));
  const float s1 = 1.0f - s0;
  const float q = 10.0f * s0 * s1;
  const float qp = 10.0f * q * (1.0f - 2.0f * s0);
  const float qL = q * kLapScale, qpL = qp * kLapScale;   // the quadratic sources of the (scaled) second-order columns
  float ub = 0.f;
  if constexpr (C::LS) {
    // summed second-order columns (forward: combine_unit): lmax = [q Qa + A s0 ; q Qb + B s1], lmin = [-q Qa + A s1 ; -q Qb + B s0]
    const float A = u[1 + D], Bq = v[1 + D];
    const float lmax0 = gmax[1 + 2 * D], lmax1 = gmax[1 + 2 * D + 1], lmin0 = gmin[1 + 2 * D], lmin1 = gmin[1 + 2 * D + 1];
    const float dl0 = lmax0 - lmin0, dl1 = lmax1 - lmin1;
    float qa = 0.f, qb = 0.f;
#pragma unroll
    for (int i = 0; i < D; ++i) {
      const float a = u[1 + i], b = v[1 + i];
      const float dmax0 = gmax[1 + i], dmax1 = gmax[1 + D + i], dmin0 = gmin[1 + i], dmin1 = gmin[1 + D + i];
      ub += q * ((dmax0 - dmin0) * a - (dmax1 - dmin1) * b);
      qa = fmaf(a, a, qa);
      qb = fmaf(b, b, qb);
      ou[1 + i] = dmax0 * s0 + dmin0 * s1 + 2.0f * qL * a * dl0;
      ov[1 + i] = dmax1 * s1 + dmin1 * s0 + 2.0f * qL * b * dl1;
    }
    ub += q * (dl0 * A - dl1 * Bq) + qpL * (dl0 * qa + dl1 * qb);
    ou[1 + D] = lmax0 * s0 + lmin0 * s1;
    ov[1 + D] = lmax1 * s1 + lmin1 * s0;
  } else {
#pragma unroll
  for (int i = 0; i < D; ++i) {
    const float a = u[1 + i], b = v[1 + i], A = u[1 + D + i], Bq = v[1 + D + i];
    const float dmax0 = gmax[1 + i], dmax1 = gmax[1 + D + i], dmin0 = gmin[1 + i], dmin1 = gmin[1 + D + i];
    const float lmax0 = gmax[1 + 2 * D + i], lmax1 = gmax[1 + 3 * D + i];
    const float lmin0 = gmin[1 + 2 * D + i], lmin1 = gmin[1 + 3 * D + i];
    const float dl0 = lmax0 - lmin0, dl1 = lmax1 - lmin1;
    ub += q * ((dmax0 - dmin0) * a - (dmax1 - dmin1) * b + dl0 * A - dl1 * Bq) + qpL * (dl0 * a * a + dl1 * b * b);
    ou[1 + i] = dmax0 * s0 + dmin0 * s1 + 2.0f * qL * a * dl0;
    ov[1 + i] = dmax1 * s1 + dmin1 * s0 + 2.0f * qL * b * dl1;
    ou[1 + D + i] = lmax0 * s0 + lmin0 * s1;
    ov[1 + D + i] = lmax1 * s1 + lmin1 * s0;
  }
  }
  ou[0] = gmax[0] * s0 + gmin[0] * s1 + ub;
  ov[0] = gmax[0] * s1 + gmin[0] * s0 - ub;
}

// Step 5: reads the adjoints of the combine output (features lane and 128+lane) from the accumulator and
// the stashed encoder outputs Y5, writes Ybar5 (layer 5 has no activation).  Returns the bias partial sum.
template <class C>
__device__ __forceinline__ float epi_combine_adjoint(const EpiPos& p, int in_col, uint32_t y_addr, uint32_t gt_addr) {
  constexpr int CGP = C::CGP, ST = C::STRIDE, EO = C::EOFF;
  float y[32];
  ld_rows32(y_addr, p, p.lane, y);
  float db = 0.f;
  if constexpr (C::PACK) {
    // packed rows (stride 10, three pairs per quarter): the adjoints of a pair are read as 16 accumulator columns
    // starting at min(10 u, 16), the results are assembled for the whole quarter and stored at once
    float oq[32];
#pragma unroll
    for (int c = 0; c < 32; ++c) oq[c] = 0.f;
#pragma unroll
    for (int u = 0; u < C::UPT; ++u) {
      constexpr int kMaxStart = 16;
      const int start = (u * ST < kMaxStart) ? u * ST : kMaxStart;
      const int off = u * ST - start;
      float gmax[16], gmin[16];
      tmem_ld<16>(p.tmem + in_col + 32 * p.rq + start, gmax);
      tmem_ld<16>(p.tmem + in_col + 128 + 32 * p.rq + start, gmin);
      combine_adjoint_unit<C>(gmax + off, gmin + off, y + u * ST, oq + u * ST);
      db += oq[u * ST] + oq[u * ST + EO];
    }
    st_rows32(gt_addr, p, p.lane, oq);
  } else {
    const uint32_t line = act_line_addr(gt_addr, p.lane, p.nb);
#pragma unroll
    for (int u = 0; u < 32 / CGP; ++u) {
      float gmax[CGP], gmin[CGP], o[CGP];
      tmem_ld<CGP>(p.tmem + in_col + 32 * p.rq + u * CGP, gmax);
      tmem_ld<CGP>(p.tmem + in_col + 128 + 32 * p.rq + u * CGP, gmin);
      combine_adjoint_unit<C>(gmax, gmin, y + u * CGP, o);
      db += o[0] + o[EO];
#pragma unroll
      for (int j = 0; j < CGP / 8; ++j) sts_chunk(line, p.lane & 7, p.cb + u * (CGP / 8) + j, o + 8 * j);
    }
  }
  return db;
}

template <int S>
struct TcRevCtl {
  uint64_t full[S], empty[S];
  uint64_t xr[2];     // Ybar operand-tile half complete (= stash-buffer half consumed)
  uint64_t ar[2];     // product half complete (= operand-tile half free again)
  uint64_t yb[2];     // stash-buffer half loaded
  uint64_t sd[2];     // stash store of operand-tile half h has finished reading shared memory
  uint32_t tmem_base;
};

constexpr int kRevRing = 5;
constexpr int kNumBiasSlots = 2432;   // sum of M_l over l = 0..12: offsets c_db[l]
static __constant__ int c_db[kNumLayers] = {0, 128, 256, 384, 512, 640, 768, 1024, 1280, 1536, 1792, 2048, 2304, 2432};

template <int D, int ORDER>
__device__ __forceinline__ void tc_reverse_role(const TcRevArgs& b, const int cta, const int ncta) {
  using C = TcCfg<D, ORDER>;
  static_assert(ORDER == 2 || ORDER == kOrderLapSum, "the loss needs second-order columns");
  constexpr int S = kRevRing;
  constexpr int CGP = C::CGP, P = C::P, NG = C::N1;
  constexpr int ACC_A = 0, ACC_B = 256;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* gt = smem;                         // Ybar operand tile
  uint8_t* xbuf = smem + kTileBytes;          // stashed forward outputs of the current layer (two halves)
  uint8_t* ring = smem + 2 * kTileBytes;
  float* sdb = reinterpret_cast<float*>(ring + S * kStageBytes);   // bias-gradient partial sums [2432] + dW13 [128] + db13
  __shared__ TcRevCtl<S> ctl;
  __shared__ float g13s[kRows];
  const TcFwdArgs& a = b.f;

  const int tid = threadIdx.x, warp = tid >> 5, lane_id = tid & 31;
  for (int i = tid; i < kNumBiasSlots + 129; i += kTcThreads) sdb[i] = 0.f;
  if (tid < kRows) g13s[tid] = 0.f;   // rows no pair owns (packed layouts) must stay zero
  if (tid == 0) {
    for (int i = 0; i < S; ++i) { mbar_init(&ctl.full[i], 1); mbar_init(&ctl.empty[i], 1); }
    for (int h = 0; h < 2; ++h) {
      mbar_init(&ctl.xr[h], kEpiThreads);
      mbar_init(&ctl.ar[h], 1);
      mbar_init(&ctl.yb[h], 1);
      mbar_init(&ctl.sd[h], 1);
    }
    fence_barrier_init();
  }
  if (warp == kMmaWarp) tmem_alloc(&ctl.tmem_base, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = ctl.tmem_base;
  const int64_t ntiles = (a.pair_count + P - 1) / P;

  if (warp == kProducerWarp) {
    if (lane_id < kProducerLanes) {
      uint32_t it = 0;
      for (int64_t tile = cta; tile < ntiles; tile += ncta) {
        for (int g = 0; g < kRevStages; ++g, ++it) {
          if (int(it % kProducerLanes) != lane_id) continue;   // the lanes take the stages in turn
          const uint32_t slot = it % S;
          ring_wait_free<S>(&ctl, it);
          mbar_arrive_expect_tx(&ctl.full[slot], kStageBytes);
          bulk_g2s(ring + slot * kStageBytes, b.rimg + size_t(g) * kStageBytes, kStageBytes, &ctl.full[slot]);
        }
      }
    }
  } else if (warp == kMmaWarp) {
    // converged warp, elected lane issues
    uint32_t it = 0, nx0 = 0, nx1 = 0;
    const uint32_t ring_addr = smem_u32(ring), gt_addr = smem_u32(gt);
    for (int64_t tile = cta; tile < ntiles; tile += ncta) {
#pragma unroll 1
      for (int l = 12; l >= 1; --l) {
        const int MH = halves(c_M[l]), OH = c_K[l] >> 7;
        const bool skip = c_rev_skip[l] != 0;
        const uint32_t acc = tmem_base + (c_rev_acc[l] ? ACC_B : ACC_A);
#pragma unroll 1
        for (int kh = 0; kh < MH; ++kh) {
          if (kh == 0) { mbar_spin(&ctl.xr[0], nx0 & 1); ++nx0; }
          else { mbar_spin(&ctl.xr[1], nx1 & 1); ++nx1; }
          tc_fence_after();
          trace_evt(a.trace, lane_id == 0 && cta == 0 && tile == int64_t(cta) + 2 * ncta, 0, l, kh);
#pragma unroll 1
          for (int h = 0; h < OH; ++h) {
            mma_block<S>(&ctl, ring_addr, gt_addr, acc + h * 128, kh, skip || kh > 0, it);
            if (kh == MH - 1) trace_evt(a.trace, lane_id == 0 && cta == 0 && tile == int64_t(cta) + 2 * ncta, 0, l, 2 + h);
            if (kh == MH - 1) {
              if (elect_one()) {
                if (MH == 1 && OH == 2) {
                  // both output halves contract over operand half 0: release it after the last block only
                  if (h == 1) {
                    umma_commit(&ctl.ar[0]);
                    umma_commit(&ctl.ar[1]);
                  }
                } else {
                  umma_commit(&ctl.ar[h]);
                }
              }
              __syncwarp();
            }
          }
        }
      }
      mbar_spin(&ctl.xr[0], nx0 & 1);   // step 0 has no product
      ++nx0;
    }
  } else if (warp == kIoWarp) {
    // ------------------------------------------------------------- stash I/O: Ybar stores, prefetch of the forward outputs
    if (lane_id == 0) {
      uint32_t nx0 = 0, nx1 = 0;
      // stash source of step l: the forward outputs X_{l+1} (Y5 for the combine step)
      auto src = [&](int l, int64_t tile) -> const uint8_t* {
        return (l == 5) ? a.st.Y5 + size_t(tile) * kHalfTileBytes
                        : a.st.X[l + 1] + size_t(tile) * size_t(c_K[l + 1]) * 256;
      };
      // Stash-buffer halves: step 12 and the generator steps use half h for features [128h, 128h+128); the
      // 128-feature encoder steps alternate (5 -> half 0, 4 -> 1, 3 -> 0, 2 -> 1, 1 -> 0, 0 -> 1) so that their
      // inputs can be fetched two steps ahead.  `load(l, tile, src_half, buf_half)`.
      auto load = [&](int l, int64_t tile, int sh, int bh) {
        if (tile >= ntiles) return;
        mbar_arrive_expect_tx(&ctl.yb[bh], kHalfTileBytes);
        stash_g2s(xbuf + bh * kHalfTileBytes, src(l, tile) + size_t(sh) * kHalfTileBytes, kHalfTileBytes, &ctl.yb[bh]);
      };
      // fused backward: flag (tile, l) is published once the bulk group holding the last half of Ybar_l has completed;
      // the check lags two groups behind the newest one, so it never waits for a store that was just issued
      uint32_t ngrp = 0, pend_g[4] = {0, 0, 0, 0}, pend_i[4] = {0, 0, 0, 0}, ph = 0, pt = 0;
      auto flush_flags = [&](uint32_t done_upto) {
        while (ph != pt && pend_g[ph & 3] <= done_upto) {
          flag_publish(b.flags + pend_i[ph & 3], b.epoch);
          ++ph;
        }
      };
      load(12, cta, 0, 0);
      load(11, cta, 1, 1);
      for (int64_t tile = cta; tile < ntiles; tile += ncta) {
        const int64_t tn = tile + ncta;
#pragma unroll 1
        for (int l = 12; l >= 0; --l) {
          const int M = c_M[l], MH = halves(M);
#pragma unroll 1
          for (int kh = 0; kh < MH; ++kh) {
            if (kh == 0) { mbar_wait(&ctl.xr[0], nx0 & 1); ++nx0; }
            else { mbar_wait(&ctl.xr[1], nx1 & 1); ++nx1; }
            // the epilogue is done with a stash-buffer half: refill it first (critical path), then store Ybar
            if (l == 12) load(11, tile, 0, 0);
            else if (l >= 7) load(l - 1, tile, kh, kh);
            else if (l == 6) { if (kh == 0) load(5, tile, 0, 0); else load(4, tile, 0, 1); }
            else if (l >= 2) load(l - 2, tile, 0, (l + 1) & 1);
            else if (l == 1) load(12, tn, 0, 0);
            else load(11, tn, 1, 1);
            stash_s2g(a.st.G[l] + size_t(tile) * M * 256 + size_t(kh) * kHalfTileBytes, gt + kh * kHalfTileBytes,
                     kHalfTileBytes);
            bulk_commit();
            bulk_wait_read<0>();
            mbar_arrive(&ctl.sd[kh]);    // the next epilogue may overwrite this half of the operand tile
            if (b.flags != nullptr) {
              ++ngrp;
              if (kh == MH - 1) {
                pend_g[pt & 3] = ngrp;
                pend_i[pt & 3] = uint32_t(tile) * kFlagStride + uint32_t(l);
                ++pt;
              }
              if (ngrp > 2) {
                bulk_wait<2>();
                flush_flags(ngrp - 2);
              }
            }
          }
        }
      }
      bulk_wait0();
      if (b.flags != nullptr) flush_flags(ngrp);
    }
  } else {
    const EpiPos p = epi_pos(tmem_base);
    const uint32_t gt_addr = smem_u32(gt), x_addr = smem_u32(xbuf), sdb_addr = smem_u32(sdb);
    const float* prm = a.params;
    const float w13f = __ldg(prm + c_w[13] + p.lane);
    long long* const detb = b.detb != nullptr ? b.detb + size_t(cta) * kDetSlots : nullptr;
    float dw13 = 0.f, unused = 0.f;
    uint32_t nar0 = 0, nar1 = 0, ny0 = 0, ny1 = 0, nsd0 = 0, nsd1 = 0;
    // The inputs of a tile's loss seeds (coordinates, output-layer rows, observed speeds of this thread's pair) reach shared
    // memory one tile ahead of their use, the gather index two tiles ahead, by asynchronous copies that occupy no registers:
    // every epilogue thread waits for the seeds at the named barrier below, so a global round trip here is a gap in every
    // tile.  Neutral for the 1 Mi-pair step, which is limited by the board's power budget (profiles/r02v_tail_loads.txt); kept
    // for the cases that leave power to spare, where the sweep is bound by this chain.
    constexpr int kSeedIn = 2 * D + C::CG + 2;
    __shared__ float seed_s[2][P][kSeedIn];
    __shared__ int32_t srow_s[2][P];
    auto pair_of = [&](int64_t tile) -> int64_t {     // this thread's pair in `tile`, or -1
      const int64_t pr = a.pair_begin + tile * P + tid;
      return (tile < ntiles && pr < a.pair_begin + a.pair_count && pr < a.n_total) ? pr : -1;
    };
    auto fetch_seed = [&](int64_t tile, int64_t pr, int64_t row, int buf) {
      if (pr < 0) return;
      float* dst = seed_s[buf][tid];
#pragma unroll
      for (int i = 0; i < 2 * D; ++i) cp_async4(dst + i, a.coords + row * a.cstride + i);
#pragma unroll
      for (int c = 0; c < C::CG; ++c) cp_async4(dst + 2 * D + c, a.st.y13 + size_t(tile) * kRows + C::pair_row(tid) + c);
      cp_async4(dst + 2 * D + C::CG, a.speed + row * a.sstride);
      cp_async4(dst + 2 * D + C::CG + 1, a.speed + row * a.sstride + 1);
    };
    if (tid < P) {
      const int64_t p0 = pair_of(cta), p1 = pair_of(int64_t(cta) + ncta);
      fetch_seed(cta, p0, p0 >= 0 ? src_row(a, p0) : 0, 0);
      if (a.index != nullptr && p1 >= 0) cp_async4(&srow_s[1][tid], a.index + p1);
    }
    int spar = 0;
    for (int64_t tile = cta; tile < ntiles; tile += ncta, spar ^= 1) {
      const int64_t pair0 = a.pair_begin + tile * P;
      int tl = 12;
      const bool tron = tid == 0 && cta == 0 && tile == int64_t(cta) + 2 * ncta;
      // ---- loss seeds and output-activation adjoint (one thread per pair), S:890-936 / M:914-951
      if (tid < P) {
        const int64_t pair = pair0 + tid;
        const bool valid = pair < a.pair_begin + a.pair_count && pair < a.n_total;
        float gy[CGP];
#pragma unroll
        for (int c = 0; c < CGP; ++c) gy[c] = 0.f;
        cp_async_wait_all();     // this tile's seed inputs and the next tile's gather index (issued a tile ago)
        if (valid) {
          float x[2 * D], y[CGP];
          const float* sin_ = seed_s[spar][tid];
#pragma unroll
          for (int i = 0; i < 2 * D; ++i) x[i] = sin_[i];
#pragma unroll
          for (int c = 0; c < C::CG; ++c) y[c] = sin_[2 * D + c];
          y[0] += __ldg(prm + c_b[13]);
          const ActOut o = actout_eval(y[0]);
          float dtau[NG], ltau[NG];
#pragma unroll
          for (int i = 0; i < NG; ++i) {
            dtau[i] = y[1 + i] * o.o1;
            ltau[i] = 0.f;
            if constexpr (!C::LS) ltau[i] = fmaf(y[1 + i] * y[1 + i], o.o2, y[1 + NG + i] * kLapUnscale * o.o1);
          }
          float qs[2] = {0.f, 0.f};
          if constexpr (C::LS) {   // per-endpoint Laplacian sums in the first slot of each endpoint (see tc_forward tail)
#pragma unroll
            for (int k = 0; k < 2; ++k) {
#pragma unroll
              for (int i = 0; i < D; ++i) qs[k] = fmaf(y[1 + k * D + i], y[1 + k * D + i], qs[k]);
              ltau[k * D] = fmaf(qs[k], o.o2, y[1 + NG + k] * kLapUnscale * o.o1);
            }
          }
          const float al = sched_alpha(a);
          const float y0 = al * sin_[2 * D + C::CG] + 1.0f - al;
          const float y1 = al * sin_[2 * D + C::CG + 1] + 1.0f - al;
          float tb, dbar[NG], lbar[NG];
          eikonal_residual<D, true>(x, o.o, dtau, ltau, y0, y1, sched_gamma(a), a.single_variant, kGradScale, &tb, dbar, lbar);
          float yb = tb * o.o1;
          if constexpr (C::LS) {
#pragma unroll
            for (int k = 0; k < 2; ++k) {
              const float lb = lbar[k * D];          // d loss / d lap_k (the same for every coordinate of endpoint k)
              yb = fmaf(lb, fmaf(qs[k], o.o3, y[1 + NG + k] * kLapUnscale * o.o2), yb);
              gy[1 + NG + k] = lb * o.o1 * kLapUnscale;     // adjoint of the scaled column
#pragma unroll
              for (int i = 0; i < D; ++i) {
                const float dy = y[1 + k * D + i];
                yb = fmaf(dbar[k * D + i] * dy, o.o2, yb);
                gy[1 + k * D + i] = fmaf(dbar[k * D + i], o.o1, lb * 2.0f * dy * o.o2);
              }
            }
          } else {
#pragma unroll
          for (int i = 0; i < NG; ++i) {
            const float dy = y[1 + i], ly = y[1 + NG + i];
            yb = fmaf(dbar[i] * dy, o.o2, yb);
            yb = fmaf(lbar[i], fmaf(dy * dy, o.o3, ly * kLapUnscale * o.o2), yb);
            gy[1 + i] = fmaf(dbar[i], o.o1, lbar[i] * 2.0f * dy * o.o2);
            gy[1 + NG + i] = lbar[i] * o.o1 * kLapUnscale;
          }
          }
          gy[0] = yb;
          if (detb != nullptr) det_add(detb + kNumBiasSlots + 128, yb);
          else atomicAdd(sdb + kNumBiasSlots + 128, yb);
        }
#pragma unroll
        for (int c = 0; c < C::STRIDE; ++c) g13s[C::pair_row(tid) + c] = gy[c];
        // the next tile's seed inputs, and the gather index of the tile after it
        const int64_t tn = tile + ncta, pn = pair_of(tn), pn2 = pair_of(tile + 2 * int64_t(ncta));
        fetch_seed(tn, pn, (a.index != nullptr && pn >= 0) ? int64_t(srow_s[spar ^ 1][tid]) : pn, spar ^ 1);
        if (a.index != nullptr && pn2 >= 0) cp_async4(&srow_s[spar][tid], a.index + pn2);
      }
      named_bar_sync(1, kEpiThreads);
      // product half ready (adjoint input, and the product is done reading operand-tile half h), the stash store of
      // that operand-tile half is done, and the stashed forward outputs of this step are loaded
#define PNTF_WAIT_SD0() { mbar_wait(&ctl.sd[0], nsd0 & 1); ++nsd0; }
#define PNTF_WAIT_SD1() { mbar_wait(&ctl.sd[1], nsd1 & 1); ++nsd1; }
#define PNTF_WAIT0_Y1() { mbar_wait(&ctl.ar[0], nar0 & 1); ++nar0; tc_fence_after(); --tl; trace_evt(a.trace, tron, 1, tl, 0); PNTF_WAIT_SD0(); mbar_wait(&ctl.yb[1], ny1 & 1); ++ny1; trace_evt(a.trace, tron, 1, tl, 1); }
#define PNTF_WAIT0() { mbar_wait(&ctl.ar[0], nar0 & 1); ++nar0; tc_fence_after(); --tl; trace_evt(a.trace, tron, 1, tl, 0); PNTF_WAIT_SD0(); mbar_wait(&ctl.yb[0], ny0 & 1); ++ny0; trace_evt(a.trace, tron, 1, tl, 1); }
      // (operand-tile half 1 was last stored by step l+1, or for l = 11 by step 6 of the previous tile)
#define PNTF_WAIT1() { mbar_wait(&ctl.ar[1], nar1 & 1); ++nar1; tc_fence_after(); if (l != 11 || tile != int64_t(cta)) PNTF_WAIT_SD1(); mbar_wait(&ctl.yb[1], ny1 & 1); ++ny1; }
#define PNTF_DB(l, h, v) { const float dbv = (v); if (detb != nullptr) det_add(detb + c_db[l] + (h) * 128 + p.lane, dbv); \
                           else red_shared_add(sdb_addr + 4u * uint32_t(c_db[l] + (h) * 128 + p.lane), dbv); }
      // ---- generator (step 12 has no accumulator input; the operand tile was last stored by step 0 of the previous tile)
      if (tile != int64_t(cta)) PNTF_WAIT_SD0();
      { mbar_wait(&ctl.yb[0], ny0 & 1); ++ny0; trace_evt(a.trace, tron, 1, 12, 1); }
      PNTF_DB(12, 0, (epi_adjoint_half<C, false, false, true>(p, 0, 0, p.lane, x_addr, gt_addr, g13s, w13f, dw13)));
      { epi_publish_each(&ctl.xr[0]); trace_evt(a.trace, tron, 1, tl, 2); }
#pragma unroll 1
      for (int l = 11; l >= 6; --l) {
        const bool second = (l & 1) != 0;      // second layer of a block: reads B, pre-loads the skip adjoint in place
        if (second) {
          PNTF_WAIT0();
          PNTF_DB(l, 0, (epi_adjoint_half<C, false, true, false>(p, ACC_B, ACC_B, p.lane, x_addr, gt_addr, nullptr, 0.f, unused)));
          { epi_publish_each(&ctl.xr[0]); trace_evt(a.trace, tron, 1, tl, 2); }
          PNTF_WAIT1();
          PNTF_DB(l, 1, (epi_adjoint_half<C, false, true, false>(p, ACC_B + 128, ACC_B + 128, 128 + p.lane, x_addr, gt_addr,
                                                                  nullptr, 0.f, unused)));
          { epi_publish_each(&ctl.xr[1]); trace_evt(a.trace, tron, 1, tl, 3); }
        } else {
          PNTF_WAIT0();
          PNTF_DB(l, 0, (epi_adjoint_half<C, false, false, false>(p, ACC_A, 0, p.lane, x_addr, gt_addr, nullptr, 0.f, unused)));
          { epi_publish_each(&ctl.xr[0]); trace_evt(a.trace, tron, 1, tl, 2); }
          PNTF_WAIT1();
          PNTF_DB(l, 1, (epi_adjoint_half<C, false, false, false>(p, ACC_A + 128, 0, 128 + p.lane, x_addr, gt_addr, nullptr,
                                                                   0.f, unused)));
          { epi_publish_each(&ctl.xr[1]); trace_evt(a.trace, tron, 1, tl, 3); }
        }
      }
      // ---- combine (needs both halves of the adjoint of its output) and encoder
      PNTF_WAIT0();
      { mbar_wait(&ctl.ar[1], nar1 & 1); ++nar1; tc_fence_after(); }
      PNTF_DB(5, 0, epi_combine_adjoint<C>(p, ACC_B, x_addr, gt_addr));
      { epi_publish_each(&ctl.xr[0]); trace_evt(a.trace, tron, 1, tl, 2); }
      PNTF_WAIT0_Y1();
      PNTF_DB(4, 0, (epi_adjoint_half<C, true, true, false>(p, ACC_A, ACC_B, p.lane, x_addr + kHalfTileBytes, gt_addr, nullptr, 0.f, unused)));
      { epi_publish_each(&ctl.xr[0]); trace_evt(a.trace, tron, 1, tl, 2); }
      PNTF_WAIT0();
      PNTF_DB(3, 0, (epi_adjoint_half<C, true, false, false>(p, ACC_A, 0, p.lane, x_addr, gt_addr, nullptr, 0.f, unused)));
      { epi_publish_each(&ctl.xr[0]); trace_evt(a.trace, tron, 1, tl, 2); }
      PNTF_WAIT0_Y1();
      PNTF_DB(2, 0, (epi_adjoint_half<C, true, true, false>(p, ACC_B, ACC_B, p.lane, x_addr + kHalfTileBytes, gt_addr, nullptr, 0.f, unused)));
      { epi_publish_each(&ctl.xr[0]); trace_evt(a.trace, tron, 1, tl, 2); }
      PNTF_WAIT0();
      PNTF_DB(1, 0, (epi_adjoint_half<C, true, false, false>(p, ACC_A, 0, p.lane, x_addr, gt_addr, nullptr, 0.f, unused)));
      { epi_publish_each(&ctl.xr[0]); trace_evt(a.trace, tron, 1, tl, 2); }
      PNTF_WAIT0_Y1();
      PNTF_DB(0, 0, (epi_adjoint_half<C, true, false, false>(p, ACC_B, 0, p.lane, x_addr + kHalfTileBytes, gt_addr, nullptr, 0.f, unused)));
      { epi_publish_each(&ctl.xr[0]); trace_evt(a.trace, tron, 1, tl, 2); }
#undef PNTF_WAIT0
#undef PNTF_WAIT0_Y1
#undef PNTF_WAIT1
#undef PNTF_DB
      // g13s is rewritten by the next tile's seeds only after every thread has left step 12 of this tile:
      // the named barrier of the next tile comes after this tile's last publish of all threads
    }
    if (detb != nullptr) det_add(detb + kNumBiasSlots + p.lane, dw13);
    else atomicAdd(sdb + kNumBiasSlots + p.lane, dw13);
    // ---- bias gradients (+ output layer weight gradient) of this CTA's tiles
    named_bar_sync(1, kEpiThreads);
    const float sc = a.sched != nullptr ? __ldg(a.sched + 1) * b.inv_scale : b.out_scale;
    float* gp = b.grad;
    if (detb == nullptr) {
      for (int l = 0; l <= 12; ++l) {
        const int M = c_M[l];
        for (int i = tid; i < M; i += kEpiThreads) atomicAdd(gp + c_b[l] + i, sdb[c_db[l] + i] * sc);
      }
      if (tid < 128) atomicAdd(gp + c_w[13] + tid, sdb[kNumBiasSlots + tid] * sc);
      if (tid == 128) atomicAdd(gp + c_b[13], sdb[kNumBiasSlots + 128] * sc);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kMmaWarp) tmem_dealloc(tmem_base, 512);
}

// stand-alone launch of the reverse sweep (PNTF_BWD_SPLIT=1: weight gradients by a separate tc_dw_kernel launch)
template <int D, int ORDER>
__global__ void __launch_bounds__(kTcThreads, 1) tc_reverse_kernel(const TcRevArgs b) {
  tc_reverse_role<D, ORDER>(b, int(blockIdx.x), int(gridDim.x));
}

}  // namespace tc
}  // namespace pntf
