// Small kernels around the hot path: weight packing, fused AdamW, loss reduction, batched path
// integration (SIMT version).
#pragma once
#include "simt_forward.cuh"

namespace pntf {

// Packed weight copies used by the SIMT kernels.  For group count G (= threads per point):
//   forward copy  F[l][k][posM(m)] = W_l[m][k],  posM(m) = (m % G) * (M/G) + m / G
//   backward copy R[l][m][posK(k)] = W_l[m][k],  posK(k) = (k % G) * (K/G) + k / G
// so that a thread's FT features (f = fg + G*j) are contiguous and load as float4.
__global__ void pack_weights_kernel(const float* __restrict__ params, float* __restrict__ fwd,
                                    float* __restrict__ bwd, int G) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= kPackedCount) return;
  int l = 0;
#pragma unroll 1
  for (int i = 1; i < kNumLayers; ++i)
    if (idx >= layer_desc(i).packed) l = i;
  const LayerDesc L = layer_desc(l);
  const int e = idx - L.packed;
  const int m = e / L.K, k = e % L.K;
  const float w = params[L.w + e];
  if (l == 13) {   // output layer row: kept as is in both copies
    if (fwd) fwd[idx] = w;
    if (bwd) bwd[idx] = w;
    return;
  }
  if (fwd) fwd[L.packed + k * L.M + (m % G) * (L.M / G) + m / G] = w;
  if (bwd) bwd[L.packed + m * L.K + (k % G) * (L.K / G) + k / G] = w;
}

// torch.optim.AdamW single step (decoupled decay; reference: models/model_res_sigmoid.py:954-956).
struct AdamArgs {
  float* p; const float* g; float* m; float* v; const uint8_t* mask; int64_t n;
  float decay;        // 1 - lr*wd
  float omb1, b2, omb2, eps;   // 1-beta1, beta2, 1-beta2 (rounded from double like torch's scalars)
  float step_size;    // lr / (1 - b1^t)
  float bc2_sqrt;     // sqrt(1 - b2^t)
  float grad_scale;
  // optional device-resident {lr, step}: the update constants are then derived on the device (CUDA-graph replay)
  const float* lr_step;
  double b1d, b2d, wdd;
};

__global__ void adamw_kernel(const AdamArgs a_in) {
  AdamArgs a = a_in;
  if (a.lr_step != nullptr) {
    __shared__ float consts[3];
    if (threadIdx.x == 0) {
      const double lr = double(a.lr_step[0]), step = double(a.lr_step[1]);
      consts[0] = float(1.0 - lr * a.wdd);
      consts[1] = float(lr / (1.0 - pow(a.b1d, step)));
      consts[2] = float(sqrt(1.0 - pow(a.b2d, step)));
    }
    __syncthreads();
    a.decay = consts[0]; a.step_size = consts[1]; a.bc2_sqrt = consts[2];
  }
  for (int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < a.n; i += int64_t(gridDim.x) * blockDim.x) {
    if (a.mask && a.mask[i] == 0) continue;
    const float g = a.g[i] * a.grad_scale;
    float p = a.p[i] * a.decay;
    float m = a.m[i];
    float v = a.v[i];
    m = m + (g - m) * a.omb1;
    v = fmaf(a.omb2 * g, g, v * a.b2);
    const float denom = sqrtf(v) / a.bc2_sqrt + a.eps;
    p = p - a.step_size * (m / denom);
    a.p[i] = p; a.m[i] = m; a.v[i] = v;
  }
}

// loss_out = {beta * loss_n, loss_n}; loss_n = sum(diff)/(E*N) [+ 0.01*|B|^2/(E*N) for the multi model]
// (models/model_res_sigmoid.py:930-934, models/model_res_sigmoid_multi.py:946-949).  One block,
// fixed summation order (deterministic).
__global__ void loss_finalize_kernel(const float* __restrict__ diff, int64_t n, const float* __restrict__ Bm, int nB,
                                     int multi, float beta, const float* __restrict__ sched,
                                     float* __restrict__ loss_out) {
  __shared__ double red[1024];
  double s = 0.0;
  for (int64_t i = threadIdx.x; i < n; i += blockDim.x) s += double(diff[i]);
  double sb = 0.0;
  if (multi)
    for (int i = threadIdx.x; i < nB; i += blockDim.x) sb += double(Bm[i]) * double(Bm[i]);
  red[threadIdx.x] = s + 0.01 * sb;
  __syncthreads();
  for (int o = blockDim.x / 2; o > 0; o >>= 1) {
    if (threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    const float loss_n = float(red[0] / double(n));
    loss_out[0] = (sched != nullptr ? sched[1] : beta) * loss_n;
    loss_out[1] = loss_n;
  }
}

// Batched bidirectional path integration (test/gib_plan.py:53-86, test/arm_plan.py:111-152): every
// CTA owns a tile of P queries and iterates  XP += eta * Gradient(XP)  on chip until each query's
// endpoints meet (|x1-x0| <= thr) or max_iter+1 steps were taken; waypoints stream to HBM.
struct PlanArgs {
  FwdArgs f;            // coords = start_goal, n_total = Q
  float eta, thr;
  int max_iter;
  float* waypoints;     // [Q][max_iter+2][2D]
  int32_t* nsteps;      // [Q]
  int fill_tail;        // 1: rows after the last step repeat the final position; 0: left untouched
};

template <int D>
__global__ void __launch_bounds__(kThreads, 1) simt_plan_kernel(const PlanArgs pa) {
  using C = Cfg<D, 1>;
  constexpr int P = C::P;
  extern __shared__ __align__(16) float arena[];
  __shared__ float y13s[C::RG];
  __shared__ TileState<D, P> ts;
  __shared__ int active[P];
  __shared__ int steps[P];
  __shared__ int any_active;
  const FwdArgs& a = pa.f;
  const int64_t ntiles = (a.pair_count + P - 1) / P;
  const int64_t wstride = int64_t(pa.max_iter + 2) * 2 * D;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t pair0 = a.pair_begin + tile * P;
    load_tile_state<C>(a, ts, pair0);
    __syncthreads();
    if (threadIdx.x < P) {
      const int p = threadIdx.x;
      float d2 = 0.f;
      for (int i = 0; i < D; ++i) { const float dv = ts.x[p][D + i] - ts.x[p][i]; d2 = fmaf(dv, dv, d2); }
      active[p] = ts.valid[p] && sqrtf(d2) > pa.thr;
      steps[p] = 0;
      if (ts.valid[p])
        for (int i = 0; i < 2 * D; ++i) pa.waypoints[(pair0 + p) * wstride + i] = ts.x[p][i];
    }
    __syncthreads();
    for (int it = 1; it <= pa.max_iter + 1; ++it) {
      if (threadIdx.x == 0) {
        int any = 0;
        for (int p = 0; p < P; ++p) any |= active[p];
        any_active = any;
      }
      __syncthreads();
      if (!any_active) break;
      forward_tile<C, false>(a, ts, arena, y13s, pair0);
      if (threadIdx.x < P && active[threadIdx.x]) {
        const int p = threadIdx.x;
        float tau, dtau[C::N1], ltau[1];
        output_rules<C>(y13s + p * C::CG, tau, dtau, ltau);
        float dir[2 * D];
        field_direction<D>(ts.x[p], tau, dtau, dir);
        float d2 = 0.f;
        for (int i = 0; i < 2 * D; ++i) ts.x[p][i] = fmaf(pa.eta, dir[i], ts.x[p][i]);
        for (int i = 0; i < D; ++i) { const float dv = ts.x[p][D + i] - ts.x[p][i]; d2 = fmaf(dv, dv, d2); }
        steps[p] = it;
        for (int i = 0; i < 2 * D; ++i) pa.waypoints[(pair0 + p) * wstride + int64_t(it) * 2 * D + i] = ts.x[p][i];
        if (!(sqrtf(d2) > pa.thr)) active[p] = 0;
      }
      __syncthreads();
    }
    // rows after the last step repeat the final position
    for (int p = 0; p < P; ++p) {
      if (!ts.valid[p]) continue;
      const int64_t base = (pair0 + p) * wstride;
      const int first = (steps[p] + 1) * 2 * D, last = pa.fill_tail ? (pa.max_iter + 2) * 2 * D : 0;
      for (int e = first + threadIdx.x; e < last; e += kThreads) pa.waypoints[base + e] = ts.x[p][e % (2 * D)];
    }
    if (threadIdx.x < P && ts.valid[threadIdx.x]) pa.nsteps[pair0 + threadIdx.x] = steps[threadIdx.x];
    __syncthreads();
  }
}

}  // namespace pntf
