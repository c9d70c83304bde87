// Per-pair scalar tail of the hot path: Eikonal residual (Model.Loss), its adjoint seeds, and the
// derived fields (TravelTimes / Speed / Speed2 / Gradient).
// Reference: models/model_res_sigmoid.py:890-936, 1169-1282; models/model_res_sigmoid_multi.py:914-951,
// 1173-1248.
#pragma once
#include <cuda_runtime.h>

namespace pntf {

// Fused batch gather (models/data_multi.py:15-31 and the in-loop `data[perm[lo:hi]]` of
// models/model_res_sigmoid_multi.py:1053-1064): when `index` is set, pair p of the batch is row index[p] of a
// GPU-resident dataset whose rows are [2d coordinates, 2 speeds] (cstride = sstride = 2d + 2); otherwise
// coords / speed are dense [n, 2d] / [n, 2] arrays.
template <class A>
__device__ __forceinline__ int64_t src_row(const A& a, int64_t pair) {
  return a.index != nullptr ? int64_t(__ldg(a.index + pair)) : pair;
}


// Device-resident schedule (pntf_loss_bwd_gather_sched / pntf_adamw_sched: an epoch of training steps replayed as a
// CUDA graph): when `sched` is set the kernels read {alpha, beta, gamma} from device memory instead of their launch
// arguments, so that a captured launch sees the values the host wrote for the current epoch (M:1011-1014, 1109).
template <class A>
__device__ __forceinline__ float sched_alpha(const A& a) { return a.sched != nullptr ? __ldg(a.sched) : a.alpha; }
template <class A>
__device__ __forceinline__ float sched_gamma(const A& a) { return a.sched != nullptr ? __ldg(a.sched + 2) : a.gamma; }

template <int D>
struct PairGeom {
  float Dv[D];   // x1 - x0
  float T0;      // |D|^2
  float a0, b0, a1, b1;   // |g0|^2, g0.D, |g1|^2, g1.D
  float S0, S1;
};

template <int D>
__device__ __forceinline__ PairGeom<D> pair_geom(const float* x, float tau, const float* dtau) {
  PairGeom<D> g;
  g.T0 = 0.f; g.a0 = 0.f; g.b0 = 0.f; g.a1 = 0.f; g.b1 = 0.f;
#pragma unroll
  for (int i = 0; i < D; ++i) {
    g.Dv[i] = x[D + i] - x[i];
    g.T0 = fmaf(g.Dv[i], g.Dv[i], g.T0);
    g.a0 = fmaf(dtau[i], dtau[i], g.a0);
    g.b0 = fmaf(dtau[i], g.Dv[i], g.b0);
    g.a1 = fmaf(dtau[D + i], dtau[D + i], g.a1);
    g.b1 = fmaf(dtau[D + i], g.Dv[i], g.b1);
  }
  g.S0 = g.T0 * g.a0 + 2.0f * tau * g.b0 + tau * tau;
  g.S1 = g.T0 * g.a1 - 2.0f * tau * g.b1 + tau * tau;
  return g;
}

// diff = sum_k (r_k + 1/r_k) - 4 with r_k = pred_k / Y_k (multi) or sqrt(pred_k / Y_k) (single),
// pred_k = 1 / (sqrt(S_k)/tau^2 + gamma * lap_k).  When `seeds` the adjoints w.r.t.
// (tau, dtau, ltau) scaled by `w` are also produced.
template <int D, bool SEEDS>
__device__ __forceinline__ float eikonal_residual(const float* x, float tau, const float* dtau, const float* ltau,
                                                  float y0, float y1, float gamma, int single_variant, float w,
                                                  float* tau_bar, float* dtau_bar, float* ltau_bar) {
  const PairGeom<D> g = pair_geom<D>(x, tau, dtau);
  const float t2 = tau * tau;
  float diff = -4.0f;
  float tb = 0.f;
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    const float S = k ? g.S1 : g.S0;
    const float Y = k ? y1 : y0;
    float lap = 0.f;
#pragma unroll
    for (int i = 0; i < D; ++i) lap += ltau[k * D + i];
    const float sq = sqrtf(S);
    const float u = sq / t2 + gamma * lap;
    const float q = 1.0f / (u * Y);
    float dl_dq;
    if (single_variant) {
      const float r = sqrtf(q);
      diff += r + 1.0f / r;
      dl_dq = (1.0f - 1.0f / q) / (2.0f * r);
    } else {
      diff += q + 1.0f / q;
      dl_dq = 1.0f - 1.0f / (q * q);
    }
    if constexpr (SEEDS) {
      const float ub = w * dl_dq * (-q / u);
      const float du_dS = 1.0f / (2.0f * sq * t2);
      const float sgn = k ? -1.0f : 1.0f;
      const float b = k ? g.b1 : g.b0;
      tb += ub * (du_dS * (2.0f * sgn * b + 2.0f * tau) - 2.0f * sq / (t2 * tau));
#pragma unroll
      for (int i = 0; i < D; ++i) {
        dtau_bar[k * D + i] = ub * du_dS * (2.0f * g.T0 * dtau[k * D + i] + sgn * 2.0f * tau * g.Dv[i]);
        ltau_bar[k * D + i] = ub * gamma;
      }
    }
  }
  if constexpr (SEEDS) *tau_bar = tb;
  return diff;
}

// Model.TravelTimes: |D| / tau
template <int D>
__device__ __forceinline__ float field_travel_time(const float* x, float tau) {
  float T0 = 0.f;
#pragma unroll
  for (int i = 0; i < D; ++i) { const float dv = x[D + i] - x[i]; T0 = fmaf(dv, dv, T0); }
  return sqrtf(T0) / tau;
}

// Model.Speed (goal side): tau^2 / sqrt(S1);  Speed2: 1/(sqrt(S1)/tau^2 + gamma*lap1)
template <int D>
__device__ __forceinline__ float field_speed(const float* x, float tau, const float* dtau, const float* ltau,
                                             float gamma, bool visc) {
  const PairGeom<D> g = pair_geom<D>(x, tau, dtau);
  if (!visc) return tau * tau / sqrtf(g.S1);
  float lap = 0.f;
#pragma unroll
  for (int i = 0; i < D; ++i) lap += ltau[D + i];
  return 1.0f / (sqrtf(g.S1) / (tau * tau) + gamma * lap);
}

// Model.Gradient: per endpoint v = -grad T, scaled by 1/|v|^2; out[2D] = (start side, goal side).
template <int D>
__device__ __forceinline__ void field_direction(const float* x, float tau, const float* dtau, float* out) {
  float Dv[D], T0 = 0.f;
#pragma unroll
  for (int i = 0; i < D; ++i) { Dv[i] = x[D + i] - x[i]; T0 = fmaf(Dv[i], Dv[i], T0); }
  const float T0s = sqrtf(T0);
  const float c1 = 1.0f / (T0s * tau);
  const float c2 = T0s / (tau * tau);
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    float v[D], n2 = 0.f;
#pragma unroll
    for (int i = 0; i < D; ++i) {
      const float V0 = k ? Dv[i] : -Dv[i];
      v[i] = -(c1 * V0 - c2 * dtau[k * D + i]);
      n2 = fmaf(v[i], v[i], n2);
    }
    const float inv = 1.0f / n2;
#pragma unroll
    for (int i = 0; i < D; ++i) out[k * D + i] = v[i] * inv;
  }
}

}  // namespace pntf
