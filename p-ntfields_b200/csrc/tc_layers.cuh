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
//
// Stage order inside a layer (forward and reverse images alike): the contraction is split into halves
// of 128 input features (kh); within kh come the output halves (h, 128 features each), each with its
// two 64-feature stages.  The MMA issuer can therefore start a layer as soon as the first 128 features
// of its operand tile exist, and the first output half completes before the second one is issued.
// Reference arithmetic: models/model_res_sigmoid.py:629-657 (linear_laplace / act_laplace).
#pragma once
#include <cuda_fp16.h>

#include "layout.cuh"
#include "tc_common.cuh"

namespace pntf {
namespace tc {

// ORDER: 0 value only, 1 value + first derivatives, 2 value + first + diagonal second derivatives (one column per
// coordinate), kOrderLapSum = value + first derivatives + ONE summed second-derivative column per endpoint.
// Model.Loss uses the Laplacian only as the per-endpoint sum lap_k = sum_i d2tau/dx_i^2 (S:905-915, M:930-944) and
// every forward-mode rule is linear in the second-order columns (lx_i <- dx_i^2 s' + lx_i s; linear layers; combine;
// output rule), so L <- s' sum_i dx_i^2 + s L is exact.  13 -> 9 columns per pair at d = 3 and 25 -> 15 at d = 6, where
// a pair then takes 16 instead of 32 tile rows (twice the pairs per tile).  Used by the loss kernels only:
// pntf_forward(LAPLACE) returns per-coordinate columns.
constexpr int kOrderLapSum = 3;

// The second-order columns travel through the network multiplied by kLapScale.  Their magnitude grows with training
// (measured: X_13 reaches 6.5e4 = the fp16 maximum after ~150 AdamW steps from random init, scripts/debug_train_nan.py)
// while the operand tiles are fp16.  Every forward-mode rule is linear in these columns -- L' = s' sum dy^2 + s L, linear
// layers, combine, output rule -- so carrying Lt = kLapScale * L only means multiplying the quadratic source terms
// (Fourier, dy^2 s', q a^2) by kLapScale and dividing once in the tail: exact in real arithmetic, free at run time.
// 2^-10 leaves room up to |L| = 6.7e7; a loss contribution gamma * L (gamma = 1e-3) is negligible long before Lt reaches
// the fp16 subnormals.  The adjoints of Lt are 1/kLapScale times those of L, which lifts them (they carry a factor gamma)
// out of the fp16 subnormal range on the way back.
constexpr float kLapScale = 1.0f / 1024.0f;
constexpr float kLapUnscale = 1024.0f;

template <int D_, int ORDER_>
struct TcCfg {
  static constexpr int D = D_, ORDER = ORDER_;
  static constexpr bool LS = (ORDER_ == kOrderLapSum);
  static constexpr int NL = LS ? 1 : (ORDER_ == 2 ? D_ : 0);       // second-order columns per endpoint
  static constexpr int NE = (ORDER_ >= 1 ? D_ : 0) + NL;           // derivative columns per endpoint
  static constexpr int CE = 1 + NE;
  static constexpr int CEP = CE <= 4 ? 4 : (CE <= 8 ? 8 : 16);      // padded columns per endpoint
  static constexpr int NG = 2 * NE;                                 // derivative columns per pair
  static constexpr int CG = 1 + NG;
  static constexpr int CGP = 2 * CEP;                               // padded columns per pair (>= CG)
  // Row packing inside the 32-row quarter an epilogue thread owns: UPT pairs of STRIDE rows each; in the encoder the
  // second endpoint of a pair starts EOFF rows after the first.  Power-of-two strides except for the d = 3 loss
  // kernels with summed Laplacian columns: 5 rows per endpoint / 9 per pair -> stride 10, three pairs per quarter
  // (rows 30, 31 of a quarter and row 9 of a pair stay zero), 12 instead of 8 pairs per tile.
  // Value-only forward (ORDER 0: NN.out, TravelTimes, Tau): one row per endpoint, stride 2, 16 pairs per quarter.
  static constexpr bool PACK = LS && D_ == 3;
  static constexpr int STRIDE = PACK ? 10 : (ORDER_ == 0 ? 2 : CGP);
  static constexpr int EOFF = PACK ? 5 : (ORDER_ == 0 ? 1 : CEP);
  static constexpr int UPT = PACK ? 3 : 32 / STRIDE;
  static constexpr int P = 4 * UPT;                                 // pairs per tile
  static constexpr int N1 = 2 * D;
  __host__ __device__ static constexpr int pair_row(int t) { return (t / UPT) * 32 + (t % UPT) * STRIDE; }
  static_assert(CG <= CGP, "column padding");
  static_assert(CGP == 8 || CGP == 16 || CGP == 32, "a pair's columns must divide a 32-row quarter");
  static_assert(EOFF >= CE && STRIDE >= 2 * EOFF && STRIDE >= CG && UPT * STRIDE <= 32, "row packing");
};

// Operand tiles the training stash need not hold (-DPNTF_STASH_REGEN=bits, bit 0: X_0, bit 1: X_6; OFF by default): X_0, the
// Fourier features, is a function of the coordinates and X_6, the combine output, of the stashed encoder outputs Y5, and
// both are read back by the weight-gradient kernel only (the reverse sweep takes its gates from X_{l+1}, l >= 0, and
// from Y5 at the combine).  With the switch on that kernel regenerates them in shared memory with sixteen otherwise idle
// warps (tc_dw.cuh): 135 KB of 757 KB of stash stores per tile leave the training forward and 97 KB of 1 311 KB of loads
// the weight-gradient kernel (ncu: 3.33 -> 3.08 MB of DRAM traffic per tile).  Parity-green (all GPU tests), but slower on
// the power-limited 1 Mi-pair step (profiles/r02r_variants.txt): the generators need ~2 900 clk per tile
// and make their work units the stragglers of the launch (17.0 -> 25 ms; 18 ms with a third of the SMs given to them),
// while the training forward turned out NOT to be write-bound at 1.6-1.7 GHz (no change with both stores removed).
#ifndef PNTF_STASH_REGEN
#define PNTF_STASH_REGEN 0
#endif
constexpr bool kStashRegen = PNTF_STASH_REGEN != 0;
__host__ __device__ constexpr bool stash_regen_layer(int l) {
  return ((PNTF_STASH_REGEN & 1) != 0 && l == 0) || ((PNTF_STASH_REGEN & 2) != 0 && l == 6);
}

constexpr int kRows = 128;                  // rows (UMMA N) per tile
constexpr int kTileBytes = 256 * kRows * 2; // activation tile: 256 features x 128 rows fp16
constexpr int kHalfTileBytes = kTileBytes / 2;
constexpr int kEpiWarps = 16;
constexpr int kEpiThreads = kEpiWarps * 32; // epilogue warps 0..15: lane quarter = warp % 4, row quarter = warp / 4
constexpr int kTcThreads = kEpiThreads + 96;   // + producer warp 16 + MMA warp 17 + stash I/O warp 18
constexpr int kProducerWarp = kEpiWarps, kMmaWarp = kEpiWarps + 1, kIoWarp = kEpiWarps + 2;

// per-layer run-time tables (constant memory: indexed with a run-time layer number by the issuer threads)
static __constant__ int c_K[kNumLayers] = {256, 128, 128, 128, 128, 128, 256, 256, 256, 256, 256, 256, 256, 128};
static __constant__ int c_M[kNumLayers] = {128, 128, 128, 128, 128, 128, 256, 256, 256, 256, 256, 256, 128, 1};
// forward: layers whose accumulator was pre-loaded with the residual skip (second layer of a block)
static __constant__ int c_fwd_skip[kNumLayers] = {0, 0, 1, 0, 1, 0, 0, 1, 0, 1, 0, 1, 0, 0};
// reverse: products W_l^T Ybar_l that land on a pre-loaded skip adjoint (first layer of a block)
static __constant__ int c_rev_skip[kNumLayers] = {0, 1, 0, 1, 0, 0, 1, 0, 1, 0, 1, 0, 0, 0};
// offsets of the layers' weights / biases inside the flat parameter buffer (layout.cuh, execution order)
static __constant__ int c_w[kNumLayers] = {0, 32896, 115328, 49408, 131840, 65920, 148352, 378753, 214144, 444545, 279936, 510337, 345728, 378624};
static __constant__ int c_b[kNumLayers] = {32768, 49280, 131712, 65792, 148224, 82304, 213888, 444289, 279680, 510081, 345472, 575873, 378496, 378752};

__host__ __device__ constexpr int halves(int features) { return (features + 127) / 128; }

// ---- forward weight stages: layer l = (K/128) x ceil(M/128) x 2 stages
__host__ __device__ constexpr int fwd_stages(int l) { return halves(layer_desc(l).K) * halves(layer_desc(l).M) * 2; }
__host__ __device__ constexpr int fwd_stage_begin(int l) {
  int s = 0;
  for (int i = 0; i < l; ++i) s += fwd_stages(i);
  return s;
}
constexpr int kFwdStages = fwd_stage_begin(kNumLayers);   // 68

// ---- reverse weight stages (Xbar_l = W_l^T Ybar_l for l = 12 .. 1): contraction over M_l, output K_l
__host__ __device__ constexpr int rev_stages(int l) {
  return (l >= 1 && l <= 12) ? halves(layer_desc(l).M) * halves(layer_desc(l).K) * 2 : 0;
}
__host__ __device__ constexpr int rev_stage_begin(int l) {   // processing order 12, 11, ..., 1
  int s = 0;
  for (int i = 12; i > l; --i) s += rev_stages(i);
  return s;
}
constexpr int kRevStages = rev_stage_begin(0);            // 62

// first stage of each layer inside the images (forward: consumption order 0..13; reverse: 12..1)
static __constant__ int c_fwd_begin[kNumLayers + 1] = {
    fwd_stage_begin(0), fwd_stage_begin(1), fwd_stage_begin(2), fwd_stage_begin(3), fwd_stage_begin(4),
    fwd_stage_begin(5), fwd_stage_begin(6), fwd_stage_begin(7), fwd_stage_begin(8), fwd_stage_begin(9),
    fwd_stage_begin(10), fwd_stage_begin(11), fwd_stage_begin(12), fwd_stage_begin(13), fwd_stage_begin(14)};
static __constant__ int c_rev_begin[kNumLayers] = {
    rev_stage_begin(0), rev_stage_begin(1), rev_stage_begin(2), rev_stage_begin(3), rev_stage_begin(4),
    rev_stage_begin(5), rev_stage_begin(6), rev_stage_begin(7), rev_stage_begin(8), rev_stage_begin(9),
    rev_stage_begin(10), rev_stage_begin(11), rev_stage_begin(12), 0};

constexpr size_t kFwdImgBytes = size_t(kFwdStages) * kStageBytes;
constexpr size_t kRevImgBytes = size_t(kRevStages) * kStageBytes;

// fp32 parameters -> fp16 stage images.  One thread = one 16-byte chunk (8 consecutive k) of one
// stage row.  Stage s of a layer: kh = s / (2*OH), h = (s/2) % OH, 64-feature block ks = 2*kh + s%2
// (OH = output halves).
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
      if (g >= c_fwd_begin[i]) l = i;
    const int K = c_K[l], M = c_M[l];
    const int s = g - c_fwd_begin[l];
    const int oh = halves(M);
    const int kh = s / (2 * oh), h = (s >> 1) % oh, ks = 2 * kh + (s & 1);
    const int mo = h * 128 + m;
    const float* W = params + c_w[l];
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = (mo < M) ? W[mo * K + ks * 64 + c * 8 + j] : 0.f;
  } else {
    int l = 12;
#pragma unroll 1
    for (int i = 11; i >= 1; --i)
      if (g >= c_rev_begin[i]) l = i;
    const int K = c_K[l];
    const int s = g - c_rev_begin[l];
    const int oh = halves(K);
    const int kh = s / (2 * oh), h = (s >> 1) % oh, ks = 2 * kh + (s & 1);
    const int ko = h * 128 + m;   // output index of the transposed product = input feature of layer l
    const float* W = params + c_w[l];
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = W[(ks * 64 + c * 8 + j) * K + ko];
  }
  uint4 q;
  q.x = pack_h2(v[0], v[1]); q.y = pack_h2(v[2], v[3]); q.z = pack_h2(v[4], v[5]); q.w = pack_h2(v[6], v[7]);
  uint8_t* img = rev ? rev_img : fwd_img;
  *reinterpret_cast<uint4*>(img + size_t(g) * kStageBytes + (m >> 3) * kAtomBytes + (m & 7) * 128 + (((c ^ (m & 7)) & 7) << 4)) = q;
}

}  // namespace tc
}  // namespace pntf
