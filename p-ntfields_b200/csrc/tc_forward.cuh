// tcgen05 forward kernel (tensor-core precision mode): a persistent, warp-specialised CTA carries
// tiles of P start/goal pairs through the whole network (Fourier features -> per-endpoint encoder ->
// symmetric combine -> generator -> output rules -> per-pair tail).  The 14 linear layers run as
// tcgen05.mma (fp16 operands, fp32 accumulators in TMEM); value / first-derivative / diagonal
// second-derivative columns are propagated in forward mode by the epilogue warps, which read the
// accumulators with tcgen05.ld, apply bias / residual skip / softplus rules and write the next
// layer's fp16 operand tile back to shared memory.  The residual skips stay in fp32: the block input
// is pre-loaded into the accumulator of the block's second layer with tcgen05.st.
//   warps 0..7 : epilogue (thread = TMEM lane = output feature, warps 0-3 rows 0..63, warps 4-7 rows 64..127)
//   warp  8    : weight-stage producer (cp.async.bulk L2 -> shared memory ring)
//   warp  9    : MMA issuer (+ bulk stores of the training stash)
// Reference: NN.out_laplace, models/model_res_sigmoid.py:676-820 (models/model_res_sigmoid_multi.py:710-848);
// NN.out_grad S:513-613; Model.Loss S:869-936 / M:897-951.
#pragma once
#include "math.cuh"
#include "pair_math.cuh"
#include "tc_layers.cuh"

namespace pntf {
namespace tc {

enum { TC_TAIL_NONE = 0, TC_TAIL_LOSS = 1, TC_TAIL_FIELD = 2 };

struct TcStash {
  uint8_t* X[kNumLayers];   // fp16 operand-tile images of the layer inputs, [tile][K_l * 256 B]
  uint8_t* Y[kNumLayers];   // fp16 tile images of the pre-activations (bias / skip included), layers 0..12
  uint8_t* G[kNumLayers];   // fp16 tile images of the pre-activation adjoints (reverse kernel), layers 0..12
  float* y13;               // [tile][128] fp32 pre-activation rows of the output layer
  float* g13;               // [tile][128] fp32 adjoints of the output layer's pre-activation
};

struct TcFwdArgs {
  const float* params;
  const uint8_t* wimg;      // forward stage images
  const float* B;           // [E][D][128]
  const float* coords;      // [n_total][2D]
  const float* speed;       // [n_total][2]
  int64_t n_total, n_per_env, pair_begin, pair_count;
  float* tau; float* dtau; float* ltau;
  float* field; int field_kind;
  float* diff; int single_variant;
  float alpha, gamma;
  int compat;
  TcStash st;
};

template <class C>
struct TcTile {
  float x[C::P][2 * C::D];
  int env[C::P];
  int valid[C::P];
};

// softplus_10 with its derivative gates, fast intrinsics (TC mode is operand-rounded to fp16 anyway)
struct FastAct { float a, s, sp; };
__device__ __forceinline__ FastAct fast_act(float y) {
  FastAct r;
  const float z = 10.0f * y;
  const float e = __expf(-fabsf(z));
  const float inv = __fdividef(1.0f, 1.0f + e);
  r.s = (z >= 0.f) ? inv : e * inv;
  r.a = 0.1f * (fmaxf(z, 0.f) + __logf(1.0f + e));
  r.sp = 10.0f * r.s * (1.0f - r.s);
  return r;
}
__device__ __forceinline__ float fast_sigmoid(float x) {
  const float e = __expf(-fabsf(x));
  const float inv = __fdividef(1.0f, 1.0f + e);
  return (x >= 0.f) ? inv : e * inv;
}

// hidden activation with the forward-mode rules on one unit: y[0] value, y[1..NDER] first derivatives,
// y[1+NDER..2*NDER] diagonal second derivatives (S:641-657)
template <int NDER, int ORDER>
__device__ __forceinline__ void act_unit(float* y, bool compat) {
  const FastAct a = fast_act(y[0]);
  y[0] = a.a;
  const float gate = compat ? fast_sigmoid(10.0f * a.a) : a.s;
#pragma unroll
  for (int i = 0; i < NDER; ++i) {
    const float dy = y[1 + i];
    if constexpr (ORDER >= 2) y[1 + NDER + i] = fmaf(dy * dy, a.sp, y[1 + NDER + i] * a.s);
    y[1 + i] = dy * gate;
  }
}

// write NC consecutive rows (starting at row block nb, chunk c0) of feature f as fp16
template <int NC>
__device__ __forceinline__ void st_cols(uint8_t* tile, int f, int nb, int c0, const float* v) {
#pragma unroll
  for (int j = 0; j < NC / 8; ++j) st_chunk(tile, f, nb, c0 + j, v + 8 * j);
}

// Epilogue of one hidden layer.  Accumulator columns: acc_col + h*128 + row.
//   ENC: a pair's CGP columns are two endpoint units of CEP columns (D derivative columns each),
//   otherwise one unit of CGP columns (2D derivative columns).
//   PRELOAD: also store the fp32 result to accumulator pre_col (residual skip of the next block).
//   TRAIN: also write the pre-activations to the Y staging tile.
template <class C, int M, bool ENC, bool PRELOAD, bool TRAIN>
__device__ __forceinline__ void epi_hidden(uint32_t tmem_lane_base, int acc_col, int pre_col,
                                           const float* __restrict__ bias, uint8_t* act, uint8_t* yst, bool compat) {
  constexpr int CGP = C::CGP, CEP = C::CEP;
  const int lane = threadIdx.x & 127, half = (threadIdx.x >> 7) & 1;
#pragma unroll 1
  for (int h = 0; h < M / 128; ++h) {
    const int f = h * 128 + lane;
    const float b = __ldg(bias + f);
#pragma unroll 1
    for (int pp = 0; pp < C::PH; ++pp) {
      const int row0 = half * 64 + pp * CGP;
      float v[CGP];
      tmem_ld<CGP>(tmem_lane_base + acc_col + h * 128 + row0, v);
      if constexpr (ENC) {
        v[0] += b;
        v[CEP] += b;
        if constexpr (TRAIN) st_cols<CGP>(yst, f, half, (pp * CGP) >> 3, v);
        act_unit<C::D, C::ORDER>(v, compat);
        act_unit<C::D, C::ORDER>(v + CEP, compat);
      } else {
        v[0] += b;
        if constexpr (TRAIN) st_cols<CGP>(yst, f, half, (pp * CGP) >> 3, v);
        act_unit<2 * C::D, C::ORDER>(v, false);
      }
      st_cols<CGP>(act, f, half, (pp * CGP) >> 3, v);
      if constexpr (PRELOAD) tmem_st<CGP>(tmem_lane_base + pre_col + h * 128 + row0, v);
    }
  }
  if constexpr (PRELOAD) tmem_st_wait();
}

// Epilogue of encoder layer 5 (no activation) fused with the symmetric combine (S:724-787): reads the
// two endpoint encodings of feature `lane`, writes smooth-max (feature lane) and smooth-min (feature
// 128+lane) with their derivative columns, and pre-loads both into the skip accumulator.
template <class C, bool TRAIN>
__device__ __forceinline__ void epi_combine(uint32_t tmem_lane_base, int acc_col, int pre_col,
                                            const float* __restrict__ bias, uint8_t* act, uint8_t* yst) {
  constexpr int CGP = C::CGP, CEP = C::CEP, D = C::D;
  const int lane = threadIdx.x & 127, half = (threadIdx.x >> 7) & 1;
  const float b = __ldg(bias + lane);
#pragma unroll 1
  for (int pp = 0; pp < C::PH; ++pp) {
    const int row0 = half * 64 + pp * CGP;
    float v[CGP];
    tmem_ld<CGP>(tmem_lane_base + acc_col + row0, v);
    v[0] += b;
    v[CEP] += b;
    if constexpr (TRAIN) st_cols<CGP>(yst, lane, half, (pp * CGP) >> 3, v);
    const float* u = v;
    const float* w = v + CEP;
    float omax[CGP], omin[CGP];
#pragma unroll
    for (int c = 0; c < CGP; ++c) { omax[c] = 0.f; omin[c] = 0.f; }
    const float dlt = u[0] - w[0];
    const float e = __expf(-10.0f * fabsf(dlt));
    const float tail = 0.1f * __logf(1.0f + e);
    omax[0] = fmaxf(u[0], w[0]) + tail;
    omin[0] = fminf(u[0], w[0]) - tail;
    if constexpr (C::ORDER >= 1) {
      const float inv = __fdividef(1.0f, 1.0f + e);
      const float s0 = (dlt >= 0.f) ? inv : e * inv;
      const float s1 = 1.0f - s0;
      const float q = 10.0f * s0 * s1;
#pragma unroll
      for (int i = 0; i < D; ++i) {
        const float da = u[1 + i], db = w[1 + i];
        omax[1 + i] = da * s0;
        omax[1 + D + i] = db * s1;
        omin[1 + i] = da * s1;
        omin[1 + D + i] = db * s0;
        if constexpr (C::ORDER >= 2) {
          const float la = u[1 + D + i], lb = w[1 + D + i];
          const float qa = q * da * da, qb = q * db * db;
          omax[1 + 2 * D + i] = qa + la * s0;
          omax[1 + 3 * D + i] = qb + lb * s1;
          omin[1 + 2 * D + i] = -qa + la * s1;
          omin[1 + 3 * D + i] = -qb + lb * s0;
        }
      }
    }
    st_cols<CGP>(act, lane, half, (pp * CGP) >> 3, omax);
    st_cols<CGP>(act, 128 + lane, half, (pp * CGP) >> 3, omin);
    tmem_st<CGP>(tmem_lane_base + pre_col + row0, omax);
    tmem_st<CGP>(tmem_lane_base + pre_col + 128 + row0, omin);
  }
  tmem_st_wait();
}

// Fourier features of both endpoints of every pair (S:204-210, M:199-213) into the layer-0 operand tile.
template <class C>
__device__ __forceinline__ void fourier_tile(const TcFwdArgs& a, const TcTile<C>& ts, uint8_t* act) {
  constexpr int CGP = C::CGP, CEP = C::CEP, D = C::D;
  const int k = threadIdx.x & 127, half = (threadIdx.x >> 7) & 1;
#pragma unroll 1
  for (int pp = 0; pp < C::PH; ++pp) {
    const int p = half * C::PH + pp;
    const float* Bp = a.B + int64_t(ts.env[p]) * D * kNumFreq + k;
    float w[D];
#pragma unroll
    for (int i = 0; i < D; ++i) w[i] = 6.2831855f * __ldg(Bp + i * kNumFreq);
    float vs[CGP], vc[CGP];
#pragma unroll
    for (int c = 0; c < CGP; ++c) { vs[c] = 0.f; vc[c] = 0.f; }
#pragma unroll
    for (int side = 0; side < 2; ++side) {
      float proj = 0.f;
#pragma unroll
      for (int i = 0; i < D; ++i) proj = fmaf(ts.x[p][side * D + i], w[i], proj);
      float sn, cs;
      sincosf(proj, &sn, &cs);
      float* us = vs + side * CEP;
      float* uc = vc + side * CEP;
      us[0] = sn;
      uc[0] = cs;
      if constexpr (C::ORDER >= 1) {
#pragma unroll
        for (int i = 0; i < D; ++i) {
          us[1 + i] = w[i] * cs;
          uc[1 + i] = -w[i] * sn;
          if constexpr (C::ORDER >= 2) {
            us[1 + D + i] = -w[i] * w[i] * sn;
            uc[1 + D + i] = -w[i] * w[i] * cs;
          }
        }
      }
    }
    st_cols<CGP>(act, k, half, (pp * CGP) >> 3, vs);
    st_cols<CGP>(act, 128 + k, half, (pp * CGP) >> 3, vc);
  }
}

// shared-memory control block
template <int S>
struct TcCtl {
  uint64_t full[S], empty[S];
  uint64_t tile_ready, acc_ready;
  uint32_t tmem_base;
};

// all epilogue threads: the operand tile (and TMEM pre-loads) of the next layer are complete
__device__ __forceinline__ void epi_publish(uint64_t* tile_ready) {
  fence_proxy_async();
  tc_fence_before();
  mbar_arrive(tile_ready);
}

// MMA issue of one layer: NH accumulator halves x KS stages of 64 input features.
//   acc_col: first accumulator column; accumulate_first: accumulator pre-loaded with the skip.
template <int S, class CTL>
__device__ __forceinline__ void mma_layer(CTL* ctl, uint32_t ring_addr, uint32_t act_addr, uint32_t tmem_base,
                                          int acc_col, int nh, int ks_count, bool accumulate_first, uint32_t& it) {
  const uint32_t idesc = make_idesc(128, kRows, false, true);
  for (int h = 0; h < nh; ++h) {
    for (int ks = 0; ks < ks_count; ++ks) {
      const uint32_t slot = it % S, ph = (it / S) & 1;
      mbar_wait(&ctl->full[slot], ph);
      tc_fence_after();
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const uint64_t ad = make_desc(ring_addr + slot * kStageBytes + j * 32, 16, 1024);
        const uint64_t bd = make_desc(act_addr + (ks * 4 + j) * 2 * kActSBO, kActLBO, kActSBO);
        umma_f16(tmem_base + acc_col + h * 128, ad, bd, idesc, (accumulate_first || ks > 0 || j > 0) ? 1u : 0u);
      }
      umma_commit(&ctl->empty[slot]);
      ++it;
    }
  }
}

template <int D, int ORDER, bool TRAIN, int TAIL>
__global__ void __launch_bounds__(kTcThreads, 1) tc_forward_kernel(const TcFwdArgs a) {
  using C = TcCfg<D, ORDER>;
  constexpr int S = TRAIN ? 5 : 8;            // weight ring depth
  constexpr int CGP = C::CGP, P = C::P;
  constexpr int ACC_A = 0, ACC_B = 256;       // TMEM accumulator regions (columns)
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* act = smem;
  uint8_t* yst = TRAIN ? smem + kTileBytes : nullptr;
  uint8_t* ring = smem + (TRAIN ? 2 : 1) * kTileBytes;
  __shared__ TcCtl<S> ctl;
  __shared__ TcTile<C> ts;
  __shared__ float y13s[kRows];

  const int tid = threadIdx.x, warp = tid >> 5, lane_id = tid & 31;
  if (tid == 0) {
    for (int i = 0; i < S; ++i) { mbar_init(&ctl.full[i], 1); mbar_init(&ctl.empty[i], 1); }
    mbar_init(&ctl.tile_ready, kEpiThreads);
    mbar_init(&ctl.acc_ready, 1);
    fence_barrier_init();
  }
  if (warp == 9) tmem_alloc(&ctl.tmem_base, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = ctl.tmem_base;
  const int64_t ntiles = (a.pair_count + P - 1) / P;

  if (warp == 8) {
    // ------------------------------------------------------------- weight-stage producer
    if (lane_id == 0) {
      uint32_t it = 0;
      for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        for (int g = 0; g < kFwdStages; ++g) {
          const uint32_t slot = it % S, ph = (it / S) & 1;
          mbar_wait(&ctl.empty[slot], ph ^ 1);
          mbar_arrive_expect_tx(&ctl.full[slot], kStageBytes);
          bulk_g2s(ring + slot * kStageBytes, a.wimg + size_t(g) * kStageBytes, kStageBytes, &ctl.full[slot]);
          ++it;
        }
      }
    }
  } else if (warp == 9) {
    // ------------------------------------------------------------- MMA issuer
    if (lane_id == 0) {
      uint32_t it = 0, nlayer = 0;
      const uint32_t ring_addr = smem_u32(ring), act_addr = smem_u32(act);
      for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int64_t ltile = tile;   // stash slot of this tile (tiles are local to the launch)
#pragma unroll 1
        for (int l = 0; l < kNumLayers; ++l) {
          mbar_wait(&ctl.tile_ready, nlayer & 1);
          tc_fence_after();
          if constexpr (TRAIN) {
            const LayerDesc L = layer_desc(l);
            if (l < 13) bulk_s2g(a.st.X[l] + size_t(ltile) * L.K * 256, act, uint32_t(L.K) * 256);
            if (l >= 1) {
              const int Mp = layer_desc(l - 1).M;
              bulk_s2g(a.st.Y[l - 1] + size_t(ltile) * Mp * 256, yst, uint32_t(Mp) * 256);
            }
            bulk_commit();
          }
          // accumulator region: second layers of the residual blocks accumulate onto the pre-loaded skip
          const bool skip = (l == 2 || l == 4 || l == 7 || l == 9 || l == 11);
          mma_layer<S>(&ctl, ring_addr, act_addr, tmem_base, skip ? ACC_B : ACC_A, fwd_mh(l), fwd_ks(l), skip, it);
          if constexpr (TRAIN) bulk_wait_read0();
          umma_commit(&ctl.acc_ready);
          ++nlayer;
        }
      }
      if constexpr (TRAIN) bulk_wait0();
    }
  } else {
    // ------------------------------------------------------------- epilogue warps
    const uint32_t lane_base = tmem_base + (uint32_t((warp & 3) * 32) << 16);
    const float* prm = a.params;
    uint32_t nacc = 0;
    const bool compat = a.compat != 0;
    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
      const int64_t pair0 = a.pair_begin + tile * P;
      if (tid < P) {
        const int64_t pair = pair0 + tid;
        const bool valid = pair < a.pair_begin + a.pair_count && pair < a.n_total;
        ts.valid[tid] = valid;
        ts.env[tid] = valid ? int(pair / a.n_per_env) : 0;
#pragma unroll
        for (int i = 0; i < 2 * D; ++i) ts.x[tid][i] = valid ? __ldg(a.coords + pair * 2 * D + i) : 0.f;
      }
      named_bar_sync(1, kEpiThreads);
      fourier_tile<C>(a, ts, act);
      epi_publish(&ctl.tile_ready);
#define PNTF_WAIT_ACC()                        \
  mbar_wait(&ctl.acc_ready, nacc & 1);         \
  ++nacc;                                      \
  tc_fence_after();
      // encoder: L0 ; {L1, L2(+skip)} ; {L3, L4(+skip)} ; L5 + combine
      PNTF_WAIT_ACC();
      epi_hidden<C, 128, true, true, TRAIN>(lane_base, ACC_A, ACC_B, prm + layer_desc(0).b, act, yst, compat);
      epi_publish(&ctl.tile_ready);
      PNTF_WAIT_ACC();
      epi_hidden<C, 128, true, false, TRAIN>(lane_base, ACC_A, 0, prm + layer_desc(1).b, act, yst, false);
      epi_publish(&ctl.tile_ready);
      PNTF_WAIT_ACC();
      epi_hidden<C, 128, true, true, TRAIN>(lane_base, ACC_B, ACC_B, prm + layer_desc(2).b, act, yst, false);
      epi_publish(&ctl.tile_ready);
      PNTF_WAIT_ACC();
      epi_hidden<C, 128, true, false, TRAIN>(lane_base, ACC_A, 0, prm + layer_desc(3).b, act, yst, false);
      epi_publish(&ctl.tile_ready);
      PNTF_WAIT_ACC();
      epi_hidden<C, 128, true, false, TRAIN>(lane_base, ACC_B, 0, prm + layer_desc(4).b, act, yst, false);
      epi_publish(&ctl.tile_ready);
      PNTF_WAIT_ACC();
      epi_combine<C, TRAIN>(lane_base, ACC_A, ACC_B, prm + layer_desc(5).b, act, yst);
      epi_publish(&ctl.tile_ready);
      // generator: 3 x {La, Lb(+skip)} ; L12 ; L13
#pragma unroll 1
      for (int ii = 0; ii < 3; ++ii) {
        PNTF_WAIT_ACC();
        epi_hidden<C, 256, false, false, TRAIN>(lane_base, ACC_A, 0, prm + layer_desc(6 + 2 * ii).b, act, yst, false);
        epi_publish(&ctl.tile_ready);
        PNTF_WAIT_ACC();
        if (ii < 2)
          epi_hidden<C, 256, false, true, TRAIN>(lane_base, ACC_B, ACC_B, prm + layer_desc(7 + 2 * ii).b, act, yst, false);
        else
          epi_hidden<C, 256, false, false, TRAIN>(lane_base, ACC_B, 0, prm + layer_desc(7 + 2 * ii).b, act, yst, false);
        epi_publish(&ctl.tile_ready);
      }
      PNTF_WAIT_ACC();
      epi_hidden<C, 128, false, false, TRAIN>(lane_base, ACC_A, 0, prm + layer_desc(12).b, act, yst, false);
      epi_publish(&ctl.tile_ready);
      PNTF_WAIT_ACC();
#undef PNTF_WAIT_ACC
      // output layer: lane 0 of the accumulator holds y13 for all rows
      if ((warp & 3) == 0) {
        const int half = warp >> 2;
#pragma unroll 1
        for (int pp = 0; pp < C::PH; ++pp) {
          float v[CGP];
          tmem_ld<CGP>(lane_base + ACC_A + half * 64 + pp * CGP, v);
          if (lane_id == 0) {
#pragma unroll
            for (int c = 0; c < CGP; ++c) y13s[half * 64 + pp * CGP + c] = v[c];
          }
        }
      }
      tc_fence_before();
      named_bar_sync(1, kEpiThreads);
      if (tid < P && ts.valid[tid]) {
        const int64_t pair = pair0 + tid;
        float y[CGP];
#pragma unroll
        for (int c = 0; c < CGP; ++c) y[c] = y13s[tid * CGP + c];
        y[0] += __ldg(prm + layer_desc(13).b);
        const ActOut o = actout_eval(y[0]);
        const float tau = o.o;
        float dtau[C::N1], ltau[C::N1];
#pragma unroll
        for (int i = 0; i < C::N1; ++i) {
          const float dy = y[1 + i];
          dtau[i] = dy * o.o1;
          ltau[i] = 0.f;
          if constexpr (ORDER >= 2) ltau[i] = fmaf(dy * dy, o.o2, y[1 + C::N1 + i] * o.o1);
        }
        if (a.tau) a.tau[pair] = tau;
        if (a.dtau)
          for (int i = 0; i < C::N1; ++i) a.dtau[pair * C::N1 + i] = dtau[i];
        if constexpr (ORDER >= 2) {
          if (a.ltau)
            for (int i = 0; i < C::N1; ++i) a.ltau[pair * C::N1 + i] = ltau[i];
        }
        const float* x = ts.x[tid];
        if constexpr (TAIL == TC_TAIL_LOSS && ORDER == 2) {
          const float y0 = a.alpha * __ldg(a.speed + pair * 2) + 1.0f - a.alpha;
          const float y1 = a.alpha * __ldg(a.speed + pair * 2 + 1) + 1.0f - a.alpha;
          a.diff[pair] = eikonal_residual<D, false>(x, tau, dtau, ltau, y0, y1, a.gamma, a.single_variant, 0.f, nullptr,
                                                    nullptr, nullptr);
        }
        if constexpr (TAIL == TC_TAIL_FIELD) {
          if (a.field_kind == 0) a.field[pair] = field_travel_time<D>(x, tau);
          if (a.field_kind == 1) a.field[pair] = field_speed<D>(x, tau, dtau, ltau, 0.f, false);
          if (a.field_kind == 3) {
            float out[2 * D];
            field_direction<D>(x, tau, dtau, out);
            for (int i = 0; i < 2 * D; ++i) a.field[pair * 2 * D + i] = out[i];
          }
          if constexpr (ORDER >= 2) {
            if (a.field_kind == 2) a.field[pair] = field_speed<D>(x, tau, dtau, ltau, a.gamma, true);
          }
        }
      }
      if constexpr (TRAIN) {
        if (tid < kRows) a.st.y13[size_t(tile) * kRows + tid] = y13s[tid];
      }
      // y13s / ts are rewritten only after the next tile's first named barrier
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 9) tmem_dealloc(tmem_base, 512);
}

}  // namespace tc
}  // namespace pntf
