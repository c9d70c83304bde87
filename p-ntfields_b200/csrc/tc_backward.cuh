// Fused backward launch of the tensor-core precision mode: the CTAs of ONE kernel are split into two roles.
//   CTAs [0, n_rev)        : reverse sweep (tc_reverse.cuh), tiles strided over these CTAs
//   CTAs [n_rev, gridDim)  : weight-gradient contraction (tc_dw.cuh), one (layer, 128-feature half) unit per CTA group,
//                            accumulators resident in TMEM for the whole launch
// The reverse sweep publishes a flag per (tile, layer) once Ybar_l of that tile is complete in global memory; the
// weight-gradient CTA of that layer picks the tile up right behind it, so its operand reads (Ybar_l just written, X_l
// just read by the sweep as the activated output of layer l-1) are served by L2 instead of HBM.  All CTAs of the launch
// are resident at once (the grid never exceeds the SM count, one CTA per SM), so waiting on the flags cannot dead-lock,
// and dependencies only point from lower to higher CTA roles of the same launch.
// The reference gets both results from autograd (loss.backward(), models/model_res_sigmoid.py:1071).
#pragma once
#include "tc_dw.cuh"
#include "tc_reverse.cuh"

namespace pntf {
namespace tc {

template <int D, int ORDER>
__global__ void __launch_bounds__(kTcThreads, 1) tc_backward_kernel(const TcRevArgs r, const TcDwArgs w, const int n_rev) {
  if (int(blockIdx.x) < n_rev) tc_reverse_role<D, ORDER>(r, int(blockIdx.x), n_rev);
  else tc_dw_role<D, ORDER>(w, int(blockIdx.x) - n_rev);
}

}  // namespace tc
}  // namespace pntf
