// fp32 SIMT forward kernel: one CTA carries a tile of P start/goal pairs through the whole network
// (Fourier features -> per-endpoint encoder -> symmetric combine -> generator -> output rules ->
// per-pair tail) with value / first-derivative / diagonal-second-derivative columns propagated in
// forward mode on chip.  Reference: NN.out_laplace, models/model_res_sigmoid.py:676-820
// (models/model_res_sigmoid_multi.py:710-848); NN.out_grad S:513-613; NN.out S:212-256.
#pragma once
#include "pair_math.cuh"
#include "simt_common.cuh"

namespace pntf {

enum { TAIL_NONE = 0, TAIL_LOSS = 1, TAIL_FIELD = 2, TAIL_PLAN = 3 };

struct FwdArgs {
  const float* params;     // flat fp32 parameters (biases, layer 13 row)
  const float* packed;     // packed forward weights (see pack_weights_kernel)
  const float* B;          // [E][D][128]
  const float* coords;     // [n_total][2D]
  const float* speed;      // [n_total][2] or null
  const int32_t* index;    // optional batch gather (pair_math.cuh: src_row), else null
  int cstride, sstride;    // row strides of coords / speed in floats (2D / 2 when dense)
  int64_t n_total;         // E*N pairs
  int64_t n_per_env;       // N
  int64_t pair_begin;      // first pair of this launch (chunking)
  int64_t pair_count;      // pairs in this launch
  float* tau; float* dtau; float* ltau;   // optional outputs
  float* field; int field_kind;           // TAIL_FIELD
  float* diff; int single_variant;        // TAIL_LOSS
  float alpha, gamma;
  const float* sched;      // optional device-resident {alpha, beta, gamma} (pair_math.cuh)
  int compat;              // derivative gate of the first encoder layer evaluated at softplus(y)
  Stash st;                // training stash (rows are local to this launch)
};

// Fourier features of the P endpoints of one side into F[256][RSE] (S:204-210, M:199-213).
// Tile state in shared memory: coordinates and environment index of the P pairs of the tile
// (pairs past the end of the launch are zero-filled and flagged invalid).
template <int D, int P>
struct TileState {
  float x[P][2 * D];
  int env[P];
  int valid[P];
};

template <class C>
__device__ __forceinline__ void load_tile_state(const FwdArgs& a, TileState<C::D, C::P>& ts, int64_t pair0) {
  if (threadIdx.x < C::P) {
    const int64_t pair = pair0 + threadIdx.x;
    const bool valid = pair < a.pair_begin + a.pair_count && pair < a.n_total;
    ts.valid[threadIdx.x] = valid;
    ts.env[threadIdx.x] = valid ? int(pair / a.n_per_env) : 0;
#pragma unroll
    for (int i = 0; i < 2 * C::D; ++i) ts.x[threadIdx.x][i] = valid ? __ldg(a.coords + src_row(a, pair) * a.cstride + i) : 0.f;
  }
}

template <class C, bool STASH>
__device__ __forceinline__ void fourier_stage(const FwdArgs& a, const TileState<C::D, C::P>& ts, float* F, int side,
                                              float* X0st, int64_t tile_local_pair0) {
  constexpr int D = C::D, CE = C::CE, RSE = C::RSE;
  for (int idx = threadIdx.x; idx < C::P * kNumFreq; idx += kThreads) {
    const int k = idx % kNumFreq, p = idx / kNumFreq;
    const float* Bp = a.B + int64_t(ts.env[p]) * D * kNumFreq + k;
    float w[D], proj = 0.f;
#pragma unroll
    for (int i = 0; i < D; ++i) {
      w[i] = 6.2831855f * __ldg(Bp + i * kNumFreq);
      proj = fmaf(ts.x[p][side * D + i], w[i], proj);
    }
    float sn, cs;
    sincosf(proj, &sn, &cs);
    float* f_s = F + k * RSE + p * CE;
    float* f_c = F + (kNumFreq + k) * RSE + p * CE;
    f_s[0] = sn;
    f_c[0] = cs;
    if constexpr (C::ORDER >= 1) {
#pragma unroll
      for (int i = 0; i < D; ++i) {
        f_s[1 + i] = w[i] * cs;
        f_c[1 + i] = -w[i] * sn;
        if constexpr (C::ORDER >= 2) {
          f_s[1 + D + i] = -w[i] * w[i] * sn;
          f_c[1 + D + i] = -w[i] * w[i] * cs;
        }
      }
    }
    if constexpr (STASH) {
      const int64_t row = ((tile_local_pair0 + p) * 2 + side) * CE;
#pragma unroll
      for (int c = 0; c < CE; ++c) {
        X0st[(row + c) * 256 + k] = f_s[c];
        X0st[(row + c) * 256 + kNumFreq + k] = f_c[c];
      }
    }
  }
}

// Symmetric smooth max / smooth min of the two endpoint encodings (S:724-787, M:752-813):
// f0,f1 [128][RSE] -> A[256][RSG]; derivative columns ordered start coordinates, goal coordinates.
template <class C, bool STASH>
__device__ __forceinline__ void combine_stage(const float* f0, const float* f1, float* A, float* X6st,
                                              int64_t tile_local_pair0) {
  constexpr int D = C::D, CE = C::CE, CG = C::CG, RSE = C::RSE, RSG = C::RSG, NE = C::NE;
  for (int idx = threadIdx.x; idx < C::P * 128; idx += kThreads) {
    const int f = idx % 128, p = idx / 128;
    const float* u = f0 + f * RSE + p * CE;
    const float* v = f1 + f * RSE + p * CE;
    float* omax = A + f * RSG + p * CG;
    float* omin = A + (128 + f) * RSG + p * CG;
    const float v0 = u[0], v1 = v[0];
    const float dlt = v0 - v1;
    const float tail = 0.1f * log1pf(expf(-10.0f * fabsf(dlt)));
    omax[0] = fmaxf(v0, v1) + tail;
    omin[0] = fminf(v0, v1) - tail;
    if constexpr (C::ORDER >= 1) {
      const float s0 = sigmoidf_acc(10.0f * dlt);
      const float s1 = 1.0f - s0;
      const float q = 10.0f * s0 * s1;
#pragma unroll
      for (int i = 0; i < D; ++i) {
        const float da = u[1 + i], db = v[1 + i];
        omax[1 + i] = da * s0;
        omax[1 + D + i] = db * s1;
        omin[1 + i] = da * s1;
        omin[1 + D + i] = db * s0;
        if constexpr (C::ORDER >= 2) {
          const float la = u[1 + D + i], lb = v[1 + D + i];
          const float qa = q * da * da, qb = q * db * db;
          omax[1 + 2 * D + i] = qa + la * s0;
          omax[1 + 3 * D + i] = qb + lb * s1;
          omin[1 + 2 * D + i] = -qa + la * s1;
          omin[1 + 3 * D + i] = -qb + lb * s0;
        }
      }
    }
    if constexpr (STASH) {
      const int64_t row = (tile_local_pair0 + p) * CG;
#pragma unroll
      for (int c = 0; c < CG; ++c) {
        X6st[(row + c) * 256 + f] = omax[c];
        X6st[(row + c) * 256 + 128 + f] = omin[c];
      }
    }
    (void)NE;
  }
}

// Whole-network forward for one tile.  Leaves the pre-activation columns of the output layer
// (layer 13) in y13s[RG] (shared memory).
template <class C, bool STASH>
__device__ __forceinline__ void forward_tile(const FwdArgs& a, const TileState<C::D, C::P>& ts, float* arena,
                                             float* y13s, int64_t pair0) {
  constexpr int P = C::P, O = C::ORDER, D = C::D;
  constexpr int RSE = C::RSE, RSG = C::RSG, CE = C::CE, CG = C::CG;
  const int64_t lp0 = pair0 - a.pair_begin;     // tile-local first pair (stash row origin)
  const float* prm = a.params;
  float* F = arena + C::F_OFF;
  float* ha = arena + C::HA_OFF;
  float* hb = arena + C::HB_OFF;
  const bool compat = a.compat != 0;
#pragma unroll 1
  for (int side = 0; side < 2; ++side) {
    float* fout = arena + (side ? C::F1_OFF : C::F0_OFF);
    const int64_t rb = (lp0 * 2 + side) * CE;
    constexpr int rps = 2 * CE;
    fourier_stage<C, STASH>(a, ts, F, side, a.st.X[0], lp0);
    __syncthreads();
    fwd_layer<P, 128, 256, O, D, ACT_HIDDEN, false, STASH>(a.packed + layer_desc(0).packed, prm + layer_desc(0).b, F, RSE,
                                                          ha, RSE, nullptr, 0, a.st.Y[0], a.st.X[1], rb, rps, compat);
    __syncthreads();
    fwd_layer<P, 128, 128, O, D, ACT_HIDDEN, false, STASH>(a.packed + layer_desc(1).packed, prm + layer_desc(1).b, ha, RSE,
                                                          hb, RSE, nullptr, 0, a.st.Y[1], a.st.X[2], rb, rps, false);
    __syncthreads();
    fwd_layer<P, 128, 128, O, D, ACT_HIDDEN, true, STASH>(a.packed + layer_desc(2).packed, prm + layer_desc(2).b, hb, RSE,
                                                         ha, RSE, ha, RSE, a.st.Y[2], a.st.X[3], rb, rps, false);
    __syncthreads();
    fwd_layer<P, 128, 128, O, D, ACT_HIDDEN, false, STASH>(a.packed + layer_desc(3).packed, prm + layer_desc(3).b, ha, RSE,
                                                          hb, RSE, nullptr, 0, a.st.Y[3], a.st.X[4], rb, rps, false);
    __syncthreads();
    fwd_layer<P, 128, 128, O, D, ACT_HIDDEN, true, STASH>(a.packed + layer_desc(4).packed, prm + layer_desc(4).b, hb, RSE,
                                                         ha, RSE, ha, RSE, a.st.Y[4], a.st.X[5], rb, rps, false);
    __syncthreads();
    fwd_layer<P, 128, 128, O, D, ACT_NONE, false, STASH>(a.packed + layer_desc(5).packed, prm + layer_desc(5).b, ha, RSE,
                                                        fout, RSE, nullptr, 0, a.st.Y[5], nullptr, rb, rps, false);
    __syncthreads();
  }
  float* A = arena;
  float* Bb = arena + C::BUF;
  combine_stage<C, STASH>(arena + C::F0_OFF, arena + C::F1_OFF, A, a.st.X[6], lp0);
  __syncthreads();
  const int64_t rbg = lp0 * CG;
#pragma unroll 1
  for (int ii = 0; ii < 3; ++ii) {
    const int la = 6 + 2 * ii, lb = 7 + 2 * ii;
    fwd_layer<P, 256, 256, O, 2 * D, ACT_HIDDEN, false, STASH>(a.packed + layer_desc(la).packed, prm + layer_desc(la).b, A, RSG,
                                                              Bb, RSG, nullptr, 0, a.st.Y[la], a.st.X[lb], rbg, CG, false);
    __syncthreads();
    fwd_layer<P, 256, 256, O, 2 * D, ACT_HIDDEN, true, STASH>(a.packed + layer_desc(lb).packed, prm + layer_desc(lb).b, Bb, RSG,
                                                             A, RSG, A, RSG, a.st.Y[lb], a.st.X[lb + 1], rbg, CG, false);
    __syncthreads();
  }
  fwd_layer<P, 128, 256, O, 2 * D, ACT_HIDDEN, false, STASH>(a.packed + layer_desc(12).packed, prm + layer_desc(12).b, A, RSG,
                                                            Bb, RSG, nullptr, 0, a.st.Y[12], a.st.X[13], rbg, CG, false);
  __syncthreads();
  // layer 13: 128 -> 1, one thread per (pair, column) row
  if (threadIdx.x < C::RG) {
    const int r = threadIdx.x;
    const float* w13 = prm + layer_desc(13).w;
    float acc = 0.f;
#pragma unroll 8
    for (int f = 0; f < 128; ++f) acc = fmaf(__ldg(w13 + f), Bb[f * RSG + r], acc);
    if (r % CG == 0) acc += __ldg(prm + layer_desc(13).b);
    y13s[r] = acc;
    if constexpr (STASH) a.st.Y[13][rbg + r] = acc;
  }
  __syncthreads();
}

// Output activation rules on one pair's layer-13 columns (S:659-674): y -> tau, dtau, ltau.
template <class C>
__device__ __forceinline__ void output_rules(const float* y, float& tau, float* dtau, float* ltau) {
  const ActOut o = actout_eval(y[0]);
  tau = o.o;
  if constexpr (C::ORDER >= 1) {
#pragma unroll
    for (int i = 0; i < C::N1; ++i) {
      const float dy = y[1 + i];
      dtau[i] = dy * o.o1;
      if constexpr (C::ORDER >= 2) ltau[i] = fmaf(dy * dy, o.o2, y[1 + C::N1 + i] * o.o1);
    }
  }
}

template <int D, int ORDER, bool STASH, int TAIL>
__global__ void __launch_bounds__(kThreads, 1) simt_forward_kernel(const FwdArgs a) {
  using C = Cfg<D, ORDER>;
  extern __shared__ __align__(16) float arena[];
  __shared__ float y13s[C::RG];
  __shared__ TileState<D, C::P> ts;
  const int64_t ntiles = (a.pair_count + C::P - 1) / C::P;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t pair0 = a.pair_begin + tile * C::P;
    load_tile_state<C>(a, ts, pair0);
    __syncthreads();
    forward_tile<C, STASH>(a, ts, arena, y13s, pair0);
    if (threadIdx.x < C::P) {
      const int64_t pair = pair0 + threadIdx.x;
      if (ts.valid[threadIdx.x]) {
        float tau, dtau[C::N1 > 0 ? C::N1 : 1], ltau[C::N1 > 0 ? C::N1 : 1];
        output_rules<C>(y13s + threadIdx.x * C::CG, tau, dtau, ltau);
        if (a.tau) a.tau[pair] = tau;
        if constexpr (ORDER >= 1) {
          if (a.dtau)
            for (int i = 0; i < C::N1; ++i) a.dtau[pair * C::N1 + i] = dtau[i];
        }
        if constexpr (ORDER >= 2) {
          if (a.ltau)
            for (int i = 0; i < C::N1; ++i) a.ltau[pair * C::N1 + i] = ltau[i];
        }
        const float* x = ts.x[threadIdx.x];
        if constexpr (TAIL == TAIL_LOSS && ORDER == 2) {
          const float* sp = a.speed + src_row(a, pair) * a.sstride;
          const float al = sched_alpha(a);
          const float y0 = al * __ldg(sp) + 1.0f - al;
          const float y1 = al * __ldg(sp + 1) + 1.0f - al;
          a.diff[pair] = eikonal_residual<D, false>(x, tau, dtau, ltau, y0, y1, sched_gamma(a), a.single_variant, 0.f,
                                                    nullptr, nullptr, nullptr);
        }
        if constexpr (TAIL == TAIL_FIELD) {
          if (a.field_kind == 0) a.field[pair] = field_travel_time<D>(x, tau);
          if constexpr (ORDER >= 1) {
            if (a.field_kind == 1) a.field[pair] = field_speed<D>(x, tau, dtau, ltau, 0.f, false);
            if (a.field_kind == 3) {
              float out[2 * D];
              field_direction<D>(x, tau, dtau, out);
              for (int i = 0; i < 2 * D; ++i) a.field[pair * 2 * D + i] = out[i];
            }
          }
          if constexpr (ORDER >= 2) {
            if (a.field_kind == 2) a.field[pair] = field_speed<D>(x, tau, dtau, ltau, a.gamma, true);
          }
        }
      }
    }
    __syncthreads();
  }
}

}  // namespace pntf
