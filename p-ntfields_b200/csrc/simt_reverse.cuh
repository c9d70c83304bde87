// fp32 SIMT reverse kernel: adjoint sweep of the second-order forward pass for one tile of pairs.
// Consumes the pre-activation stash written by simt_forward_kernel<.., STASH=true, ..>, seeds the
// adjoints from the Eikonal loss, walks the 14 layers backwards on chip (activation adjoints,
// dX = Ybar W, skip and combine adjoints) and writes the pre-activation adjoints Ybar_l to the G
// stash; the weight gradients dW_l = Ybar_l^T X_l are reduced afterwards by dw_kernel.
// The reference gets all of this from autograd (loss.backward(), models/model_res_sigmoid.py:1071);
// formulas: SURVEY.md appendix A.2.
#pragma once
#include "simt_forward.cuh"

namespace pntf {

struct BwdArgs {
  FwdArgs f;
  const float* packed_bwd;   // packed backward weights
  float wscale;              // d loss / d diff_n  = beta / (E*N)
  float inv_count;           // 1 / (E*N): with a device-resident schedule wscale = sched[1] * inv_count
};

// Adjoint of combine_stage.  Inputs: Gb[256][RSG] = adjoint of the combine output (max features
// 0..127, min features 128..255); the stashed encoder outputs Y5 of both sides.  Outputs: adjoints
// of the two encoder outputs, g0/g1 [128][RSE], also written to the G stash of layer 5.
template <class C>
__device__ __forceinline__ void combine_bwd(const float* Gb, const float* __restrict__ Y5, float* g0, float* g1,
                                            float* __restrict__ G5, int64_t lp0) {
  constexpr int D = C::D, CE = C::CE, CG = C::CG, RSE = C::RSE, RSG = C::RSG;
  for (int idx = threadIdx.x; idx < C::P * 128; idx += kThreads) {
    const int f = idx % 128, p = idx / 128;
    const int64_t r0 = ((lp0 + p) * 2 + 0) * CE, r1 = ((lp0 + p) * 2 + 1) * CE;
    float u[CE], v[CE];
#pragma unroll
    for (int c = 0; c < CE; ++c) { u[c] = Y5[(r0 + c) * 128 + f]; v[c] = Y5[(r1 + c) * 128 + f]; }
    const float* gmax = Gb + f * RSG + p * CG;
    const float* gmin = Gb + (128 + f) * RSG + p * CG;
    const float s0 = sigmoidf_acc(10.0f * (u[0] - v[0]));
    const float s1 = 1.0f - s0;
    const float q = 10.0f * s0 * s1;
    const float qp = 10.0f * q * (1.0f - 2.0f * s0);
    float ub = 0.f;
    float ou[CE], ov[CE];
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
#pragma unroll
    for (int c = 0; c < CE; ++c) {
      g0[f * RSE + p * CE + c] = ou[c];
      g1[f * RSE + p * CE + c] = ov[c];
      G5[(r0 + c) * 128 + f] = ou[c];
      G5[(r1 + c) * 128 + f] = ov[c];
    }
  }
}

template <int D>
__global__ void __launch_bounds__(kThreads, 1) simt_reverse_kernel(const BwdArgs b) {
  using C = Cfg<D, 2>;
  constexpr int P = C::P, CE = C::CE, CG = C::CG, RSE = C::RSE, RSG = C::RSG, NG = C::N1;
  extern __shared__ __align__(16) float arena[];
  __shared__ float g13s[C::RG];
  __shared__ TileState<D, P> ts;
  const FwdArgs& a = b.f;
  const float* pb = b.packed_bwd;
  const int64_t ntiles = (a.pair_count + P - 1) / P;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t pair0 = a.pair_begin + tile * P;
    const int64_t lp0 = pair0 - a.pair_begin;
    const int64_t rbg = lp0 * CG;
    load_tile_state<C>(a, ts, pair0);
    __syncthreads();
    // ---- loss seeds and output-activation adjoint (per pair) ----
    if (threadIdx.x < P) {
      const int p = threadIdx.x;
      const int64_t pair = pair0 + p;
      float y[CG], gy[CG];
#pragma unroll
      for (int c = 0; c < CG; ++c) y[c] = a.st.Y[13][rbg + p * CG + c];
      if (ts.valid[p]) {
        const ActOut o = actout_eval(y[0]);
        float tau = o.o, dtau[NG], ltau[NG];
#pragma unroll
        for (int i = 0; i < NG; ++i) {
          dtau[i] = y[1 + i] * o.o1;
          ltau[i] = fmaf(y[1 + i] * y[1 + i], o.o2, y[1 + NG + i] * o.o1);
        }
        const float* sp = a.speed + src_row(a, pair) * a.sstride;
        const float al = sched_alpha(a);
        const float y0 = al * __ldg(sp) + 1.0f - al;
        const float y1 = al * __ldg(sp + 1) + 1.0f - al;
        float tb, db[NG], lb[NG];
        const float diff = eikonal_residual<D, true>(ts.x[p], tau, dtau, ltau, y0, y1, sched_gamma(a), a.single_variant,
                                                     (a.sched != nullptr ? __ldg(a.sched + 1) * b.inv_count : b.wscale), &tb, db, lb);
        if (a.diff) a.diff[pair] = diff;
        float yb = tb * o.o1;
#pragma unroll
        for (int i = 0; i < NG; ++i) {
          const float dy = y[1 + i], ly = y[1 + NG + i];
          yb = fmaf(db[i] * dy, o.o2, yb);
          yb = fmaf(lb[i], fmaf(dy * dy, o.o3, ly * o.o2), yb);
          gy[1 + i] = fmaf(db[i], o.o1, lb[i] * 2.0f * dy * o.o2);
          gy[1 + NG + i] = lb[i] * o.o1;
        }
        gy[0] = yb;
      } else {
#pragma unroll
        for (int c = 0; c < CG; ++c) gy[c] = 0.f;
      }
#pragma unroll
      for (int c = 0; c < CG; ++c) {
        g13s[p * CG + c] = gy[c];
        a.st.G[13][rbg + p * CG + c] = gy[c];
      }
    }
    __syncthreads();
    float* A = arena;
    float* Bb = arena + C::BUF;
    // ---- layer 13 (128 -> 1): Xbar13[f][row] = W13[f] * Ybar13[row] ----
    {
      const float* w13 = a.params + layer_desc(13).w;
      for (int idx = threadIdx.x; idx < 128 * C::RG; idx += kThreads) {
        const int r = idx % C::RG, f = idx / C::RG;
        A[f * RSG + r] = __ldg(w13 + f) * g13s[r];
      }
    }
    __syncthreads();
    act_bwd<P, 128, NG>(A, RSG, a.st.Y[12], a.st.G[12], rbg, CG);
    __syncthreads();
    bwd_dx<P, 256, 128, CG, false>(pb + layer_desc(12).packed, A, RSG, Bb, RSG);
    __syncthreads();
#pragma unroll 1
    for (int ii = 2; ii >= 0; --ii) {
      const int la = 6 + 2 * ii, lb = 7 + 2 * ii;
      act_bwd<P, 256, NG>(Bb, RSG, a.st.Y[lb], a.st.G[lb], rbg, CG);
      __syncthreads();
      bwd_dx<P, 256, 256, CG, false>(pb + layer_desc(lb).packed, Bb, RSG, A, RSG);
      __syncthreads();
      act_bwd<P, 256, NG>(A, RSG, a.st.Y[la], a.st.G[la], rbg, CG);
      __syncthreads();
      bwd_dx<P, 256, 256, CG, true>(pb + layer_desc(la).packed, A, RSG, Bb, RSG);
      __syncthreads();
    }
    // ---- combine adjoint: Bb -> g0, g1 (placed in A's region) ----
    float* g0 = arena;
    float* g1 = arena + 128 * RSE;
    combine_bwd<C>(Bb, a.st.Y[5], g0, g1, a.st.G[5], lp0);
    __syncthreads();
    float* ha = arena + C::BUF;
    float* hb = ha + 128 * RSE;
#pragma unroll 1
    for (int side = 0; side < 2; ++side) {
      const float* gs = side ? g1 : g0;
      const int64_t rb = (lp0 * 2 + side) * CE;
      constexpr int rps = 2 * CE;
      bwd_dx<P, 128, 128, CE, false>(pb + layer_desc(5).packed, gs, RSE, ha, RSE);
      __syncthreads();
#pragma unroll 1
      for (int ii = 1; ii >= 0; --ii) {
        const int la = 1 + 2 * ii, lb = 2 + 2 * ii;
        act_bwd<P, 128, D>(ha, RSE, a.st.Y[lb], a.st.G[lb], rb, rps);
        __syncthreads();
        bwd_dx<P, 128, 128, CE, false>(pb + layer_desc(lb).packed, ha, RSE, hb, RSE);
        __syncthreads();
        act_bwd<P, 128, D>(hb, RSE, a.st.Y[la], a.st.G[la], rb, rps);
        __syncthreads();
        bwd_dx<P, 128, 128, CE, true>(pb + layer_desc(la).packed, hb, RSE, ha, RSE);
        __syncthreads();
      }
      act_bwd<P, 128, D>(ha, RSE, a.st.Y[0], a.st.G[0], rb, rps);
      __syncthreads();
    }
  }
}

// dW[m][k] += sum_rows G[row][m] * X[row][k]   (+ bias gradient: rows whose column index is 0)
// One CTA = a 128x128 output tile over one slice of rows; split-K partial sums are added to
// grad_params with atomics.
struct DwArgs {
  const float* G; const float* X;
  float* dW; float* db;
  int M, K;
  int64_t rows;
  int cols_per_point;     // rows come in groups of this many columns; column 0 feeds the bias
  int64_t rows_per_split;
};

__global__ void __launch_bounds__(256) dw_kernel(const DwArgs d) {
  constexpr int TM = 128, TK = 128, TR = 16;
  __shared__ __align__(16) float Gs[TR][TM];
  __shared__ __align__(16) float Xs[TR][TK];
  const int m0 = blockIdx.x * TM, k0 = blockIdx.y * TK;
  const int64_t r_begin = int64_t(blockIdx.z) * d.rows_per_split;
  int64_t r_end = r_begin + d.rows_per_split;
  if (r_end > d.rows) r_end = d.rows;
  const int tm = threadIdx.x / 16, tk = threadIdx.x % 16;   // 16 x 16 threads, 8 x 8 outputs each
  float acc[8][8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
  float bacc = 0.f;   // bias partial: threads 0..127 own feature m0 + threadIdx.x (k-tile 0 only)
  for (int64_t r0 = r_begin; r0 < r_end; r0 += TR) {
    for (int idx = threadIdx.x; idx < TR * TM / 4; idx += 256) {
      const int r = idx / (TM / 4), c4 = idx % (TM / 4);
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (r0 + r < r_end) {
        if (d.M >= TM) {
          v = __ldg(reinterpret_cast<const float4*>(d.G + (r0 + r) * d.M + m0) + c4);
        } else {   // M == 1 (output layer)
          if (c4 == 0) v.x = __ldg(d.G + (r0 + r));
        }
      }
      *reinterpret_cast<float4*>(&Gs[r][c4 * 4]) = v;
    }
    for (int idx = threadIdx.x; idx < TR * TK / 4; idx += 256) {
      const int r = idx / (TK / 4), c4 = idx % (TK / 4);
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (r0 + r < r_end) v = __ldg(reinterpret_cast<const float4*>(d.X + (r0 + r) * d.K + k0) + c4);
      *reinterpret_cast<float4*>(&Xs[r][c4 * 4]) = v;
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < TR; ++r) {
      float g[8], x[8];
      const float4 g0 = *reinterpret_cast<const float4*>(&Gs[r][tm * 4]);
      const float4 g1 = *reinterpret_cast<const float4*>(&Gs[r][64 + tm * 4]);
      const float4 x0 = *reinterpret_cast<const float4*>(&Xs[r][tk * 4]);
      const float4 x1 = *reinterpret_cast<const float4*>(&Xs[r][64 + tk * 4]);
      g[0] = g0.x; g[1] = g0.y; g[2] = g0.z; g[3] = g0.w; g[4] = g1.x; g[5] = g1.y; g[6] = g1.z; g[7] = g1.w;
      x[0] = x0.x; x[1] = x0.y; x[2] = x0.z; x[3] = x0.w; x[4] = x1.x; x[5] = x1.y; x[6] = x1.z; x[7] = x1.w;
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(g[i], x[j], acc[i][j]);
    }
    if (blockIdx.y == 0 && threadIdx.x < TM) {
      for (int r = 0; r < TR; ++r)
        if (r0 + r < r_end && (r0 + r) % d.cols_per_point == 0) bacc += Gs[r][threadIdx.x];
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int m = m0 + (i < 4 ? tm * 4 + i : 64 + tm * 4 + (i - 4));
    if (m >= d.M) continue;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int k = k0 + (j < 4 ? tk * 4 + j : 64 + tk * 4 + (j - 4));
      atomicAdd(d.dW + size_t(m) * d.K + k, acc[i][j]);
    }
  }
  if (blockIdx.y == 0 && threadIdx.x < TM && m0 + threadIdx.x < d.M) atomicAdd(d.db + m0 + threadIdx.x, bacc);
}

}  // namespace pntf
