// fp32 SIMT ("verify mode") building blocks: tile configuration, the column GEMM every layer uses,
// the activation epilogue and its adjoint.  One CTA of 256 threads owns a tile of P start/goal
// pairs; activations live in shared memory as X[feature][row] with row = point*C + column
// (column 0 = value, then the first-derivative columns, then the diagonal second-derivative
// columns) so that one thread owns FT features of one point and sees all of its columns.
//
// Reference arithmetic: models/model_res_sigmoid.py:629-657 (linear_laplace / act_laplace).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "layout.cuh"
#include "math.cuh"

namespace pntf {

constexpr int kThreads = 256;

template <int D_, int ORDER_>
struct Cfg {
  static constexpr int D = D_, ORDER = ORDER_;
  static constexpr int P = (D <= 3) ? 8 : 4;       // pairs per tile
  static constexpr int G = kThreads / P;            // feature groups (threads per point)
  static constexpr int NE = ORDER * D, CE = 1 + NE; // derivative / total columns per endpoint
  static constexpr int NG = 2 * NE, CG = 1 + NG;    // ... per pair in the generator
  static constexpr int N1 = (ORDER >= 1) ? 2 * D : 0;  // first-derivative columns per pair (= size of dtau / ltau rows)
  static constexpr int RE = P * CE, RSE = RE | 1;   // rows per encoder pass, odd smem row stride
  static constexpr int RG = P * CG, RSG = RG | 1;
  static constexpr int BUF = 256 * RSG;             // one generator activation buffer (floats)
  // encoder-pass regions inside the arena
  static constexpr int F_OFF = 0;                   // Fourier features [256][RSE]
  static constexpr int HA_OFF = 256 * RSE;          // hidden ping [128][RSE]
  static constexpr int HB_OFF = HA_OFF + 128 * RSE; // hidden pong
  static constexpr int ENC_END = HB_OFF + 128 * RSE;
  static constexpr int F0_OFF = (ENC_END > BUF) ? ENC_END : BUF;  // encoder output, start side
  static constexpr int F1_OFF = F0_OFF + 128 * RSE;               // encoder output, goal side
  static constexpr int ARENA = (F1_OFF + 128 * RSE > 2 * BUF) ? (F1_OFF + 128 * RSE) : 2 * BUF;
  static constexpr size_t SMEM_BYTES = size_t(ARENA) * sizeof(float) + 1024;
};

// Per-layer global stash pointers for training (fp32, [row][feature]).
struct Stash {
  float* X[kNumLayers];   // layer inputs
  float* Y[kNumLayers];   // pre-activations (bias and skip included)
  float* G[kNumLayers];   // adjoints of the pre-activations (written by the reverse kernel)
};

// acc[j][c] += sum_k Wp[k][fg*FT + j] * Xs[k][p*NC + c]
// Wp is a packed weight copy: [K][M] with feature f stored at (f % G)*FT + f / G.
template <int P, int MOUT, int KIN, int NC>
__device__ __forceinline__ void gemm_cols(const float* __restrict__ Wp, const float* __restrict__ Xs, int xs,
                                          float (&acc)[MOUT / (kThreads / P)][NC]) {
  constexpr int G = kThreads / P;
  constexpr int FT = MOUT / G;
  static_assert(FT == 2 || FT == 4 || FT == 8, "feature tile");
  const int fg = threadIdx.x % G;
  const int p = threadIdx.x / G;
  const float* wp = Wp + fg * FT;
  const float* xp = Xs + p * NC;
#pragma unroll 2
  for (int k = 0; k < KIN; ++k) {
    float w[FT];
    if constexpr (FT == 8) {
      const float4 a = __ldg(reinterpret_cast<const float4*>(wp + size_t(k) * MOUT));
      const float4 b = __ldg(reinterpret_cast<const float4*>(wp + size_t(k) * MOUT) + 1);
      w[0] = a.x; w[1] = a.y; w[2] = a.z; w[3] = a.w; w[4] = b.x; w[5] = b.y; w[6] = b.z; w[7] = b.w;
    } else if constexpr (FT == 4) {
      const float4 a = __ldg(reinterpret_cast<const float4*>(wp + size_t(k) * MOUT));
      w[0] = a.x; w[1] = a.y; w[2] = a.z; w[3] = a.w;
    } else {
      const float2 a = __ldg(reinterpret_cast<const float2*>(wp + size_t(k) * MOUT));
      w[0] = a.x; w[1] = a.y;
    }
    float x[NC];
#pragma unroll
    for (int c = 0; c < NC; ++c) x[c] = xp[k * xs + c];
#pragma unroll
    for (int j = 0; j < FT; ++j)
#pragma unroll
      for (int c = 0; c < NC; ++c) acc[j][c] = fmaf(w[j], x[c], acc[j][c]);
  }
}

enum { ACT_NONE = 0, ACT_HIDDEN = 1 };

// One forward layer on a tile: Y = X W^T (+b on the value column) (+skip); optional stash of Y;
// hidden activation with derivative rules; result to Ys (and to the next layer's X stash).
//   NDER = number of first-derivative columns; ORDER picks how many column groups exist.
template <int P, int MOUT, int KIN, int ORDER, int NDER, int ACT, bool SKIP, bool STASH>
__device__ __forceinline__ void fwd_layer(const float* __restrict__ Wp, const float* __restrict__ bias,
                                          const float* Xs, int xs, float* Ys, int ys, const float* Sk, int ss,
                                          float* __restrict__ Yst, float* __restrict__ Xnext_st,
                                          int64_t row_base, int row_pstride, bool compat_gate) {
  constexpr int G = kThreads / P;
  constexpr int FT = MOUT / G;
  constexpr int NC = 1 + ORDER * NDER;
  const int fg = threadIdx.x % G;
  const int p = threadIdx.x / G;
  float acc[FT][NC];
#pragma unroll
  for (int j = 0; j < FT; ++j)
#pragma unroll
    for (int c = 0; c < NC; ++c) acc[j][c] = 0.0f;
  gemm_cols<P, MOUT, KIN, NC>(Wp, Xs, xs, acc);
  const int64_t row = row_base + int64_t(p) * row_pstride;
#pragma unroll
  for (int j = 0; j < FT; ++j) {
    const int f = fg + G * j;
    float* y = acc[j];
    y[0] += __ldg(bias + f);
    if constexpr (SKIP) {
#pragma unroll
      for (int c = 0; c < NC; ++c) y[c] += Sk[f * ss + p * NC + c];
    }
    if constexpr (STASH) {
#pragma unroll
      for (int c = 0; c < NC; ++c) Yst[(row + c) * MOUT + f] = y[c];
    }
    if constexpr (ACT == ACT_HIDDEN) {
      const Act a = act_eval(y[0]);
      y[0] = a.a;
      if constexpr (ORDER >= 1) {
        const float gate = compat_gate ? sigmoidf_acc(10.0f * a.a) : a.s;
#pragma unroll
        for (int i = 0; i < NDER; ++i) {
          const float dy = y[1 + i];
          if constexpr (ORDER >= 2) y[1 + NDER + i] = fmaf(dy * dy, a.sp, y[1 + NDER + i] * a.s);
          y[1 + i] = dy * gate;
        }
      }
    }
#pragma unroll
    for (int c = 0; c < NC; ++c) Ys[f * ys + p * NC + c] = y[c];
    if constexpr (STASH) {
      if (Xnext_st != nullptr) {
#pragma unroll
        for (int c = 0; c < NC; ++c) Xnext_st[(row + c) * MOUT + f] = y[c];
      }
    }
  }
}

// Adjoint of the hidden activation, in place on Gs[f][row] (ORDER 2 only).  Reads the stashed
// pre-activations, writes the pre-activation adjoints back to Gs and to the G stash.
//   x_bar  = g0 s + sum_i [ gd_i dy_i s' + gl_i (dy_i^2 s'' + ly_i s') ]
//   dy_bar = gd_i s + gl_i 2 dy_i s' ;  ly_bar = gl_i s            
template <int P, int M, int NDER>
__device__ __forceinline__ void act_bwd(float* Gs, int gs, const float* __restrict__ Yst, float* __restrict__ Gst,
                                        int64_t row_base, int row_pstride) {
  constexpr int G = kThreads / P;
  constexpr int FT = M / G;
  constexpr int NC = 1 + 2 * NDER;
  const int fg = threadIdx.x % G;
  const int p = threadIdx.x / G;
  const int64_t row = row_base + int64_t(p) * row_pstride;
#pragma unroll 1
  for (int j = 0; j < FT; ++j) {
    const int f = fg + G * j;
    float y[NC], g[NC];
#pragma unroll
    for (int c = 0; c < NC; ++c) y[c] = Yst[(row + c) * M + f];
#pragma unroll
    for (int c = 0; c < NC; ++c) g[c] = Gs[f * gs + p * NC + c];
    const Act a = act_eval(y[0]);
    const float spp = act_spp(a);
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
#pragma unroll
    for (int c = 0; c < NC; ++c) {
      Gs[f * gs + p * NC + c] = g[c];
      Gst[(row + c) * M + f] = g[c];
    }
  }
}

// Xbar[k][row] (+)= sum_f Wb[f][k] * Ybar[f][row]  (adjoint of a linear layer w.r.t. its input).
// Wb is the packed "backward" copy: [M][K] with input feature k stored at (k % G)*FT + k / G.
template <int P, int KOUT, int MRED, int NC, bool ACCUM>
__device__ __forceinline__ void bwd_dx(const float* __restrict__ Wb, const float* Gs, int gs, float* Xb, int xbs) {
  constexpr int G = kThreads / P;
  constexpr int FT = KOUT / G;
  const int fg = threadIdx.x % G;
  const int p = threadIdx.x / G;
  float acc[FT][NC];
#pragma unroll
  for (int j = 0; j < FT; ++j)
#pragma unroll
    for (int c = 0; c < NC; ++c) acc[j][c] = ACCUM ? Xb[(fg + G * j) * xbs + p * NC + c] : 0.0f;
  gemm_cols<P, KOUT, MRED, NC>(Wb, Gs, gs, acc);
#pragma unroll
  for (int j = 0; j < FT; ++j)
#pragma unroll
    for (int c = 0; c < NC; ++c) Xb[(fg + G * j) * xbs + p * NC + c] = acc[j][c];
}

}  // namespace pntf
