// tcgen05 reverse kernel (tensor-core precision mode): adjoint sweep of the second-order forward pass.
// Same CTA organisation as tc_forward_kernel.  Per tile it seeds the adjoints from the Eikonal loss,
// then walks the layers backwards: the epilogue warps read Xbar_{l+1} (adjoint of layer l's activated
// output) from TMEM and the stashed pre-activations Y_l from shared memory, apply the activation
// adjoint, write Ybar_l as the fp16 operand tile of the next product
// Xbar_l = W_l^T Ybar_l (tcgen05.mma, transposed weight stages) and pre-load it in fp32 into the TMEM
// accumulator that receives the residual-skip adjoint.  Every Ybar_l tile is also dumped to the stash
// (over Y_l) for the weight-gradient contraction (tc_dw_kernel); bias gradients and the 128 -> 1 output
// layer's weight gradient are reduced in registers.
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
};

// read NC consecutive rows of feature f (fp16 -> fp32)
template <int NC>
__device__ __forceinline__ void ld_cols(const uint8_t* tile, int f, int nb, int c0, float* v) {
#pragma unroll
  for (int j = 0; j < NC / 8; ++j) {
    const uint4 q = *reinterpret_cast<const uint4*>(tile + act_chunk_off(f, nb, c0 + j));
    const __half2* h = reinterpret_cast<const __half2*>(&q);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 t = __half22float2(h[i]);
      v[8 * j + 2 * i] = t.x;
      v[8 * j + 2 * i + 1] = t.y;
    }
  }
}

// adjoint of act_unit, in place on g (incoming: adjoints of the activated outputs), y = pre-activations
template <int NDER>
__device__ __forceinline__ void act_adjoint_unit(float* g, const float* y) {
  const FastAct a = fast_act(y[0]);
  const float spp = 10.0f * a.sp * (1.0f - 2.0f * a.s);
  float xb = g[0] * a.s;
#pragma unroll
  for (int i = 0; i < NDER; ++i) {
    const float dy = y[1 + i], ly = y[1 + NDER + i];
    const float gd = g[1 + i], gl = g[1 + NDER + i];
    xb = fmaf(gd * dy, a.sp, xb);
    xb = fmaf(gl, fmaf(dy * dy, spp, ly * a.sp), xb);
    g[1 + i] = fmaf(gd, a.s, gl * 2.0f * dy * a.sp);
    g[1 + NDER + i] = gl * a.s;
  }
  g[0] = xb;
}

// Ybar_l for one hidden layer.  SRC13: the incoming adjoint is the rank-1 product w13[f] * g13[row]
// (output layer) instead of a TMEM accumulator; its weight gradient is reduced here as well.
template <class C, int M, bool ENC, bool PRELOAD, bool SRC13>
__device__ __forceinline__ void epi_adjoint(uint32_t tmem_lane_base, int acc_col, int pre_col, const uint8_t* ybuf,
                                            uint8_t* gt, float* db, const float* g13s, float w13f, float& dw13) {
  constexpr int CGP = C::CGP, CEP = C::CEP;
  const int lane = threadIdx.x & 127, half = (threadIdx.x >> 7) & 1;
#pragma unroll
  for (int h = 0; h < M / 128; ++h) {
    const int f = h * 128 + lane;
    float dbh = 0.f;
#pragma unroll 1
    for (int pp = 0; pp < C::PH; ++pp) {
      const int row0 = half * 64 + pp * CGP;
      float g[CGP], y[CGP];
      if constexpr (SRC13) {
#pragma unroll
        for (int c = 0; c < CGP; ++c) g[c] = w13f * g13s[row0 + c];
      } else {
        tmem_ld<CGP>(tmem_lane_base + acc_col + h * 128 + row0, g);
      }
      ld_cols<CGP>(ybuf, f, half, (pp * CGP) >> 3, y);
      if constexpr (SRC13) {
        // X13 = act(Y12): weight gradient of the output layer, dW13[f] += sum_c g13[c] X13[f][c]
        float x[CGP];
#pragma unroll
        for (int c = 0; c < CGP; ++c) x[c] = y[c];
        act_unit<2 * C::D, C::ORDER>(x, false);
#pragma unroll
        for (int c = 0; c < C::CG; ++c) dw13 = fmaf(g13s[row0 + c], x[c], dw13);
      }
      if constexpr (ENC) {
        act_adjoint_unit<C::D>(g, y);
        act_adjoint_unit<C::D>(g + CEP, y + CEP);
        dbh += g[0] + g[CEP];
      } else {
        act_adjoint_unit<2 * C::D>(g, y);
        dbh += g[0];
      }
      st_cols<CGP>(gt, f, half, (pp * CGP) >> 3, g);
      if constexpr (PRELOAD) tmem_st<CGP>(tmem_lane_base + pre_col + h * 128 + row0, g);
    }
    db[h] += dbh;
  }
  if constexpr (PRELOAD) tmem_st_wait();
}

// Adjoint of the symmetric combine (S:724-787): reads the adjoints of the
// smooth-max (feature lane) and smooth-min (feature 128+lane) outputs from the accumulator and the stashed
// encoder outputs Y5 of both endpoints, writes Ybar5 (= adjoint of the encoder output, layer 5 has no
// activation) for both endpoints.
template <class C>
__device__ __forceinline__ void epi_combine_adjoint(uint32_t tmem_lane_base, int acc_col, const uint8_t* ybuf,
                                                    uint8_t* gt, float* db) {
  constexpr int CGP = C::CGP, CEP = C::CEP, D = C::D;
  const int lane = threadIdx.x & 127, half = (threadIdx.x >> 7) & 1;
  float dbh = 0.f;
#pragma unroll 1
  for (int pp = 0; pp < C::PH; ++pp) {
    const int row0 = half * 64 + pp * CGP;
    float gmax[CGP], gmin[CGP], y[CGP], o[CGP];
    tmem_ld<CGP>(tmem_lane_base + acc_col + row0, gmax);
    tmem_ld<CGP>(tmem_lane_base + acc_col + 128 + row0, gmin);
    ld_cols<CGP>(ybuf, lane, half, (pp * CGP) >> 3, y);
#pragma unroll
    for (int c = 0; c < CGP; ++c) o[c] = 0.f;
    const float* u = y;
    const float* v = y + CEP;
    float* ou = o;
    float* ov = o + CEP;
    const float s0 = fast_sigmoid(10.0f * (u[0] - v[0]));
    const float s1 = 1.0f - s0;
    const float q = 10.0f * s0 * s1;
    const float qp = 10.0f * q * (1.0f - 2.0f * s0);
    float ub = 0.f;
#pragma unroll
    for (int i = 0; i < D; ++i) {
      const float a = u[1 + i], b = v[1 + i], A = u[1 + D + i], Bq = v[1 + D + i];
      const float dmax0 = gmax[1 + i], dmax1 = gmax[1 + D + i], dmin0 = gmin[1 + i], dmin1 = gmin[1 + D + i];
      const float lmax0 = gmax[1 + 2 * D + i], lmax1 = gmax[1 + 3 * D + i];
      const float lmin0 = gmin[1 + 2 * D + i], lmin1 = gmin[1 + 3 * D + i];
      const float dl0 = lmax0 - lmin0, dl1 = lmax1 - lmin1;
      ub += q * ((dmax0 - dmin0) * a - (dmax1 - dmin1) * b + dl0 * A - dl1 * Bq) + qp * (dl0 * a * a + dl1 * b * b);
      ou[1 + i] = dmax0 * s0 + dmin0 * s1 + 2.0f * q * a * dl0;
      ov[1 + i] = dmax1 * s1 + dmin1 * s0 + 2.0f * q * b * dl1;
      ou[1 + D + i] = lmax0 * s0 + lmin0 * s1;
      ov[1 + D + i] = lmax1 * s1 + lmin1 * s0;
    }
    ou[0] = gmax[0] * s0 + gmin[0] * s1 + ub;
    ov[0] = gmax[0] * s1 + gmin[0] * s0 - ub;
    dbh += ou[0] + ov[0];
    st_cols<CGP>(gt, lane, half, (pp * CGP) >> 3, o);
  }
  db[0] += dbh;
}

template <int S>
struct TcRevCtl {
  uint64_t full[S], empty[S];
  uint64_t tile_ready, acc_ready, y_full;
  uint32_t tmem_base;
};

// accumulate-onto-skip layers of the reverse sweep (their product lands on the pre-loaded skip adjoint)
__host__ __device__ constexpr bool rev_on_skip(int l) { return l == 10 || l == 8 || l == 6 || l == 3 || l == 1; }

template <int D>
__global__ void __launch_bounds__(kTcThreads, 1) tc_reverse_kernel(const TcRevArgs b) {
  using C = TcCfg<D, 2>;
  constexpr int S = 5;
  constexpr int CGP = C::CGP, P = C::P, NG = C::N1;
  constexpr int ACC_A = 0, ACC_B = 256;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* gt = smem;                         // Ybar operand tile
  uint8_t* ybuf = smem + kTileBytes;          // stashed pre-activations of the current layer
  uint8_t* ring = smem + 2 * kTileBytes;
  __shared__ TcRevCtl<S> ctl;
  __shared__ TcTile<C> ts;
  __shared__ float g13s[kRows];
  const TcFwdArgs& a = b.f;

  const int tid = threadIdx.x, warp = tid >> 5, lane_id = tid & 31;
  if (tid == 0) {
    for (int i = 0; i < S; ++i) { mbar_init(&ctl.full[i], 1); mbar_init(&ctl.empty[i], 1); }
    mbar_init(&ctl.tile_ready, kEpiThreads);
    mbar_init(&ctl.acc_ready, 1);
    mbar_init(&ctl.y_full, 1);
    fence_barrier_init();
  }
  if (warp == 9) tmem_alloc(&ctl.tmem_base, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = ctl.tmem_base;
  const int64_t ntiles = (a.pair_count + P - 1) / P;

  if (warp == 8) {
    if (lane_id == 0) {
      uint32_t it = 0;
      for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        for (int g = 0; g < kRevStages; ++g) {
          const uint32_t slot = it % S, ph = (it / S) & 1;
          mbar_wait(&ctl.empty[slot], ph ^ 1);
          mbar_arrive_expect_tx(&ctl.full[slot], kStageBytes);
          bulk_g2s(ring + slot * kStageBytes, b.rimg + size_t(g) * kStageBytes, kStageBytes, &ctl.full[slot]);
          ++it;
        }
      }
    }
  } else if (warp == 9) {
    if (lane_id == 0) {
      uint32_t it = 0, n = 0;
      const uint32_t ring_addr = smem_u32(ring), gt_addr = smem_u32(gt);
      mbar_arrive(&ctl.acc_ready);     // completion #0: the operand tile starts out free
      if (int64_t(blockIdx.x) < ntiles) {
        mbar_arrive_expect_tx(&ctl.y_full, 128 * 256);
        bulk_g2s(ybuf, a.st.Y[12] + size_t(blockIdx.x) * 128 * 256, 128 * 256, &ctl.y_full);
      }
      for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
#pragma unroll 1
        for (int l = 12; l >= 0; --l) {
          mbar_wait(&ctl.tile_ready, n & 1);
          tc_fence_after();
          const int Ml = layer_desc(l).M;
          bulk_s2g(a.st.G[l] + size_t(tile) * Ml * 256, gt, uint32_t(Ml) * 256);
          bulk_commit();
          // pre-activations of the next step (all epilogue threads are done with the current ones)
          const int ln = (l > 0) ? l - 1 : 12;
          const int64_t tn = (l > 0) ? tile : tile + gridDim.x;
          if (tn < ntiles) {
            const uint32_t bytes = uint32_t(layer_desc(ln).M) * 256;
            mbar_arrive_expect_tx(&ctl.y_full, bytes);
            bulk_g2s(ybuf, a.st.Y[ln] + size_t(tn) * bytes, bytes, &ctl.y_full);
          }
          if (l >= 1)
            mma_layer<S>(&ctl, ring_addr, gt_addr, tmem_base, rev_on_skip(l) ? ACC_B : ACC_A, rev_mh(l), rev_ks(l),
                         rev_on_skip(l), it);
          bulk_wait_read0();
          if (l >= 1) umma_commit(&ctl.acc_ready);
          else mbar_arrive(&ctl.acc_ready);
          ++n;
        }
      }
      bulk_wait0();
    }
  } else {
    const uint32_t lane_base = tmem_base + (uint32_t((warp & 3) * 32) << 16);
    const float* prm = a.params;
    const int lane = tid & 127;
    const float w13f = __ldg(prm + layer_desc(13).w + lane);
    float dbs[19];
#pragma unroll
    for (int i = 0; i < 19; ++i) dbs[i] = 0.f;
    float dw13 = 0.f, db13 = 0.f, unused = 0.f;
    uint32_t n = 0;
    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
      const int64_t pair0 = a.pair_begin + tile * P;
      // ---- loss seeds and output-activation adjoint (one thread per pair), S:890-936 / M:914-951
      if (tid < P) {
        const int64_t pair = pair0 + tid;
        const bool valid = pair < a.pair_begin + a.pair_count && pair < a.n_total;
        float gy[CGP];
#pragma unroll
        for (int c = 0; c < CGP; ++c) gy[c] = 0.f;
        if (valid) {
          float x[2 * D], y[CGP];
#pragma unroll
          for (int i = 0; i < 2 * D; ++i) x[i] = __ldg(a.coords + pair * 2 * D + i);
#pragma unroll
          for (int c = 0; c < CGP; ++c) y[c] = a.st.y13[size_t(tile) * kRows + tid * CGP + c];
          y[0] += __ldg(prm + layer_desc(13).b);
          const ActOut o = actout_eval(y[0]);
          float dtau[NG], ltau[NG];
#pragma unroll
          for (int i = 0; i < NG; ++i) {
            dtau[i] = y[1 + i] * o.o1;
            ltau[i] = fmaf(y[1 + i] * y[1 + i], o.o2, y[1 + NG + i] * o.o1);
          }
          const float y0 = a.alpha * __ldg(a.speed + pair * 2) + 1.0f - a.alpha;
          const float y1 = a.alpha * __ldg(a.speed + pair * 2 + 1) + 1.0f - a.alpha;
          float tb, dbar[NG], lbar[NG];
          eikonal_residual<D, true>(x, o.o, dtau, ltau, y0, y1, a.gamma, a.single_variant, kGradScale, &tb, dbar, lbar);
          float yb = tb * o.o1;
#pragma unroll
          for (int i = 0; i < NG; ++i) {
            const float dy = y[1 + i], ly = y[1 + NG + i];
            yb = fmaf(dbar[i] * dy, o.o2, yb);
            yb = fmaf(lbar[i], fmaf(dy * dy, o.o3, ly * o.o2), yb);
            gy[1 + i] = fmaf(dbar[i], o.o1, lbar[i] * 2.0f * dy * o.o2);
            gy[1 + NG + i] = lbar[i] * o.o1;
          }
          gy[0] = yb;
          db13 += yb;
        }
#pragma unroll
        for (int c = 0; c < CGP; ++c) g13s[tid * CGP + c] = gy[c];
      }
      named_bar_sync(1, kEpiThreads);
#define PNTF_WAIT_STEP()                     \
  mbar_wait(&ctl.acc_ready, n & 1);          \
  tc_fence_after();                          \
  mbar_wait(&ctl.y_full, n & 1);             \
  ++n;
      // generator
      PNTF_WAIT_STEP();
      epi_adjoint<C, 128, false, false, true>(lane_base, 0, 0, ybuf, gt, dbs + 0, g13s, w13f, dw13);
      epi_publish(&ctl.tile_ready);
      PNTF_WAIT_STEP();
      epi_adjoint<C, 256, false, true, false>(lane_base, ACC_A, ACC_B, ybuf, gt, dbs + 1, nullptr, 0.f, unused);
      epi_publish(&ctl.tile_ready);
      PNTF_WAIT_STEP();
      epi_adjoint<C, 256, false, false, false>(lane_base, ACC_A, 0, ybuf, gt, dbs + 3, nullptr, 0.f, unused);
      epi_publish(&ctl.tile_ready);
      PNTF_WAIT_STEP();
      epi_adjoint<C, 256, false, true, false>(lane_base, ACC_B, ACC_B, ybuf, gt, dbs + 5, nullptr, 0.f, unused);
      epi_publish(&ctl.tile_ready);
      PNTF_WAIT_STEP();
      epi_adjoint<C, 256, false, false, false>(lane_base, ACC_A, 0, ybuf, gt, dbs + 7, nullptr, 0.f, unused);
      epi_publish(&ctl.tile_ready);
      PNTF_WAIT_STEP();
      epi_adjoint<C, 256, false, true, false>(lane_base, ACC_B, ACC_B, ybuf, gt, dbs + 9, nullptr, 0.f, unused);
      epi_publish(&ctl.tile_ready);
      PNTF_WAIT_STEP();
      epi_adjoint<C, 256, false, false, false>(lane_base, ACC_A, 0, ybuf, gt, dbs + 11, nullptr, 0.f, unused);
      epi_publish(&ctl.tile_ready);
      // combine and encoder
      PNTF_WAIT_STEP();
      epi_combine_adjoint<C>(lane_base, ACC_B, ybuf, gt, dbs + 13);
      epi_publish(&ctl.tile_ready);
      PNTF_WAIT_STEP();
      epi_adjoint<C, 128, true, true, false>(lane_base, ACC_A, ACC_B, ybuf, gt, dbs + 14, nullptr, 0.f, unused);
      epi_publish(&ctl.tile_ready);
      PNTF_WAIT_STEP();
      epi_adjoint<C, 128, true, false, false>(lane_base, ACC_A, 0, ybuf, gt, dbs + 15, nullptr, 0.f, unused);
      epi_publish(&ctl.tile_ready);
      PNTF_WAIT_STEP();
      epi_adjoint<C, 128, true, true, false>(lane_base, ACC_B, ACC_B, ybuf, gt, dbs + 16, nullptr, 0.f, unused);
      epi_publish(&ctl.tile_ready);
      PNTF_WAIT_STEP();
      epi_adjoint<C, 128, true, false, false>(lane_base, ACC_A, 0, ybuf, gt, dbs + 17, nullptr, 0.f, unused);
      epi_publish(&ctl.tile_ready);
      PNTF_WAIT_STEP();
      epi_adjoint<C, 128, true, false, false>(lane_base, ACC_B, 0, ybuf, gt, dbs + 18, nullptr, 0.f, unused);
      epi_publish(&ctl.tile_ready);
#undef PNTF_WAIT_STEP
    }
    // ---- bias gradients (+ output layer weight gradient) of this CTA's tiles
    const float sc = b.out_scale;
    float* gp = b.grad;
    atomicAdd(gp + layer_desc(12).b + lane, dbs[0] * sc);
#pragma unroll
    for (int j = 0; j < 6; ++j) {            // layers 11 .. 6, two feature halves each
      const int l = 11 - j;
      atomicAdd(gp + layer_desc(l).b + lane, dbs[1 + 2 * j] * sc);
      atomicAdd(gp + layer_desc(l).b + 128 + lane, dbs[2 + 2 * j] * sc);
    }
#pragma unroll
    for (int j = 0; j < 6; ++j) atomicAdd(gp + layer_desc(5 - j).b + lane, dbs[13 + j] * sc);
    atomicAdd(gp + layer_desc(13).w + lane, dw13 * sc);
    if (tid < P) atomicAdd(gp + layer_desc(13).b, db13 * sc);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 9) tmem_dealloc(tmem_base, 512);
}

}  // namespace tc
}  // namespace pntf
