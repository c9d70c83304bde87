// Tensor-core path: tile geometry, weight-stage tables and the packing kernel.
//
// A tile is 128 "rows" = P start/goal pairs x CGP columns (value, first-derivative and diagonal
// second-derivative columns of the forward-mode propagation, padded to a power of two).  In the
// encoder a pair's CGP columns hold its two endpoints (CEP columns each); in the generator they hold
// the pair's 1 + 2*ORDER*D columns.  Activations live in shared memory as fp16 [feature][row] tiles in
// the SWIZZLE_128B MN-major operand layout of tc_common.cuh, so a layer is
//     D[out-feature (TMEM lane)][row (TMEM column)] = W[out][in] * X[in][row]
// with W streamed from L2 in 16 KB stages ([128 out] x [64 in], K-major) and the epilogue thread of
// lane f seeing all columns of a pair contiguously.
// Reference arithmetic: models/model_res_sigmoid.py:629-657 (linear_laplace / act_laplace).
#pragma once
#include <cuda_fp16.h>

#include "layout.cuh"
#include "tc_common.cuh"

namespace pntf {
namespace tc {

template <int D_, int ORDER_>
struct TcCfg {
  static constexpr int D = D_, ORDER = ORDER_;
  static constexpr int NE = ORDER * D;                              // derivative columns per endpoint
  static constexpr int CE = 1 + NE;
  static constexpr int CEP = CE <= 4 ? 4 : (CE <= 8 ? 8 : 16);      // padded columns per endpoint
  static constexpr int NG = 2 * NE;                                 // derivative columns per pair
  static constexpr int CG = 1 + NG;
  static constexpr int CGP = 2 * CEP;                               // padded columns per pair (>= CG)
  static constexpr int P = 128 / CGP;                               // pairs per tile
  static constexpr int PH = P / 2;                                  // pairs per 64-row block
  static constexpr int N1 = 2 * D;
  static_assert(CG <= CGP, "column padding");
};

constexpr int kRows = 128;                  // rows (UMMA N) per tile
constexpr int kTileBytes = 256 * kRows * 2; // activation tile: 256 features x 128 rows fp16
constexpr int kEpiThreads = 256;            // epilogue warps 0..7
constexpr int kTcThreads = 320;             // + producer warp 8 + MMA warp 9

// ---- forward weight stages: layer l = ceil(M/128) halves x K/64 stages, in consumption order
__host__ __device__ constexpr int fwd_mh(int l) { return layer_desc(l).M > 128 ? 2 : 1; }
__host__ __device__ constexpr int fwd_ks(int l) { return layer_desc(l).K / 64; }
__host__ __device__ constexpr int fwd_stages(int l) { return fwd_mh(l) * fwd_ks(l); }
__host__ __device__ constexpr int fwd_stage_begin(int l) {
  int s = 0;
  for (int i = 0; i < l; ++i) s += fwd_stages(i);
  return s;
}
constexpr int kFwdStages = fwd_stage_begin(kNumLayers);   // 68

// ---- reverse weight stages (Xbar_l = W_l^T Ybar_l for l = 12 .. 1): K_l/128 halves x M_l/64 stages
__host__ __device__ constexpr int rev_mh(int l) { return layer_desc(l).K / 128; }
__host__ __device__ constexpr int rev_ks(int l) { return layer_desc(l).M / 64; }
__host__ __device__ constexpr int rev_stages(int l) { return (l >= 1 && l <= 12) ? rev_mh(l) * rev_ks(l) : 0; }
__host__ __device__ constexpr int rev_stage_begin(int l) {   // processing order 12, 11, ..., 1
  int s = 0;
  for (int i = 12; i > l; --i) s += rev_stages(i);
  return s;
}
constexpr int kRevStages = rev_stage_begin(0);            // 62

constexpr size_t kFwdImgBytes = size_t(kFwdStages) * kStageBytes;
constexpr size_t kRevImgBytes = size_t(kRevStages) * kStageBytes;

// fp32 parameters -> fp16 stage images.  One thread = one 16-byte chunk (8 consecutive k) of one
// stage row.
__global__ void pack_images_kernel(const float* __restrict__ params, uint8_t* __restrict__ fwd_img,
                                   uint8_t* __restrict__ rev_img) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const int total_f = kFwdStages * 1024, total_r = kRevStages * 1024;
  if (idx >= total_f + total_r) return;
  const bool rev = idx >= total_f;
  if (rev && rev_img == nullptr) return;
  const int e = rev ? idx - total_f : idx;
  const int g = e >> 10, m = (e >> 3) & 127, c = e & 7;
  float v[8];
  if (!rev) {
    int l = 0;
#pragma unroll 1
    for (int i = 1; i < kNumLayers; ++i)
      if (g >= fwd_stage_begin(i)) l = i;
    const LayerDesc L = layer_desc(l);
    const int s = g - fwd_stage_begin(l);
    const int h = s / fwd_ks(l), ks = s % fwd_ks(l);
    const int mo = h * 128 + m;
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = (mo < L.M) ? params[L.w + mo * L.K + ks * 64 + c * 8 + j] : 0.f;
  } else {
    int l = 12;
#pragma unroll 1
    for (int i = 11; i >= 1; --i)
      if (g >= rev_stage_begin(i)) l = i;
    const LayerDesc L = layer_desc(l);
    const int s = g - rev_stage_begin(l);
    const int h = s / rev_ks(l), ks = s % rev_ks(l);
    const int ko = h * 128 + m;   // output index of the transposed product = input feature of layer l
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = params[L.w + (ks * 64 + c * 8 + j) * L.K + ko];
  }
  uint4 q;
  q.x = pack_h2(v[0], v[1]); q.y = pack_h2(v[2], v[3]); q.z = pack_h2(v[4], v[5]); q.w = pack_h2(v[6], v[7]);
  uint8_t* img = rev ? rev_img : fwd_img;
  *reinterpret_cast<uint4*>(img + size_t(g) * kStageBytes + (m >> 3) * kAtomBytes + (m & 7) * 128 + (((c ^ (m & 7)) & 7) << 4)) = q;
}

}  // namespace tc
}  // namespace pntf
