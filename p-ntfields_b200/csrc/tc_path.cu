// Host side of the tensor-core precision mode (PNTF_PREC_TC): workspace carving and launches of the
// tcgen05 kernels.  Called from the C-ABI entry points in capi.cu.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>

#include "../../include/pntf_b200.h"
#include "tc_api.cuh"
#include "tc_backward.cuh"

namespace pntf {

using namespace tc;

static size_t align256(size_t x) { return (x + 255) / 256 * 256; }

template <typename K>
static int tc_set_smem(K kernel, size_t bytes) {
  cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(bytes));
  if (e != cudaSuccess) return fail(PNTF_ERR_CUDA, "cudaFuncSetAttribute(smem=%zu): %s", bytes, cudaGetErrorString(e));
  return PNTF_OK;
}

static int tc_pack(const float* params, uint8_t* fwd_img, uint8_t* rev_img, cudaStream_t s) {
  const int total = (kFwdStages + (rev_img ? kRevStages : 0)) * 1024;
  prof_start(KC_PACK, s);
  pack_images_kernel<<<(total + 255) / 256, 256, 0, s>>>(params, fwd_img, rev_img);
  prof_stop(s);
  return launch_check("tc_pack_images");
}

constexpr size_t kFwdSmem = 2 * size_t(kTileBytes) + size_t(kFwdRing) * kStageBytes + 1024;
constexpr size_t kRevSmem = 2 * size_t(kTileBytes) + size_t(kRevRing) * kStageBytes + (kNumBiasSlots + 132) * sizeof(float) + 1024;

// Debug aid (PNTF_TRACE=1 in the environment): the forward kernel records clock64 time stamps of the
// hand-over events of CTA 0's third tile; they are printed to stderr after a stream synchronisation.
// Never active in normal runs (allocates and synchronises).
static long long* trace_begin() {
  static const bool on = getenv("PNTF_TRACE") != nullptr;
  if (!on) return nullptr;
  long long* buf = nullptr;
  if (cudaMalloc(&buf, 3 * 16 * 4 * sizeof(long long)) != cudaSuccess) return nullptr;
  cudaMemset(buf, 0, 3 * 16 * 4 * sizeof(long long));
  return buf;
}
static void trace_end(long long* buf, cudaStream_t s, const char* what) {
  if (buf == nullptr) return;
  long long h[3 * 16 * 4];
  cudaStreamSynchronize(s);
  cudaMemcpy(h, buf, sizeof(h), cudaMemcpyDeviceToHost);
  cudaFree(buf);
  long long t0 = 0;
  for (int i = 0; i < 3 * 16 * 4; ++i)
    if (h[i] != 0 && (t0 == 0 || h[i] < t0)) t0 = h[i];
  fprintf(stderr, "PNTF_TRACE %s (cycles since first event; role 0 = MMA issuer: operand half 0/1 seen, commit half 0/1 issued; "
                  "role 1/2 = epilogue thread 0 / last: acc half 0 ready, half 0 published, acc half 1 ready, half 1 published)\n", what);
  for (int role = 0; role < 3; ++role)
    for (int l = 0; l < 16; ++l) {
      const long long* e = h + (role * 16 + l) * 4;
      if (e[0] == 0 && e[1] == 0 && e[2] == 0 && e[3] == 0) continue;
      fprintf(stderr, "  role %d layer %2d :", role, l);
      for (int k = 0; k < 4; ++k) fprintf(stderr, " %8lld", e[k] ? e[k] - t0 : -1LL);
      fprintf(stderr, "\n");
    }
}

template <int D, int ORDER, bool TRAIN, int TAIL>
static int tc_launch_forward(const TcFwdArgs& a_in, cudaStream_t s) {
  using C = TcCfg<D, ORDER>;
  TcFwdArgs a = a_in;
  a.trace = trace_begin();
  auto k = tc_forward_kernel<D, ORDER, TRAIN, TAIL>;
  const size_t smem = kFwdSmem;
  if (int rc = tc_set_smem(k, smem)) return rc;
  const int64_t ntiles = (a.pair_count + C::P - 1) / C::P;
  const int grid = int(ntiles < sm_count() ? ntiles : sm_count());
  prof_start(KC_TC_FORWARD, s);
  k<<<grid, kTcThreads, smem, s>>>(a);
  prof_stop(s);
  trace_end(a.trace, s, "tc_forward");
  return launch_check("tc_forward");
}

// training stash bytes per tile: operand tiles X_l (l = 0..13), encoder outputs Y5, adjoint tiles Ybar_l
// (l = 0..12) and the fp32 output rows
static size_t stash_tile_bytes() {
  size_t s = 0;
  for (int l = 0; l <= 13; ++l) s += stash_regen_layer(l) ? 0 : size_t(layer_desc(l).K) * 256;   // X_0, X_6: regenerated (tc_layers.cuh)
  for (int l = 0; l <= 12; ++l) s += size_t(layer_desc(l).M) * 256;
  return s + kHalfTileBytes + kRows * sizeof(float) + kFlagStride * sizeof(uint32_t);
}
// Column mode of the loss kernels: summed Laplacian columns (TcCfg: kOrderLapSum, default) or one second-order column
// per coordinate (PNTF_LOSS_COLUMNS=full; same results, kept as a cross-check).
static int loss_order() {
  static const int order = [] {
    const char* e = getenv("PNTF_LOSS_COLUMNS");
    return (e != nullptr && strcmp(e, "full") == 0) ? 2 : kOrderLapSum;
  }();
  return order;
}
static int tc_tile_pairs(int d) {
  if (loss_order() == kOrderLapSum) return d <= 3 ? TcCfg<3, kOrderLapSum>::P : TcCfg<6, kOrderLapSum>::P;
  return d <= 3 ? TcCfg<3, 2>::P : TcCfg<6, 2>::P;
}
constexpr size_t kImgBytes = ((kFwdImgBytes + kRevImgBytes + 255) / 256) * 256;
// deterministic mode: fixed-point accumulators detw[PNTF_PARAM_COUNT] and detb[kDetMaxCtas][kDetSlots] behind the images
constexpr int kDetMaxCtas = 160;
constexpr size_t kDetWBytes = ((size_t(PNTF_PARAM_COUNT) * 8 + 255) / 256) * 256;
constexpr size_t kDetBytes = kDetWBytes + ((size_t(kDetMaxCtas) * kDetSlots * 8 + 255) / 256) * 256;
constexpr size_t kHeadBytes = kImgBytes + kDetBytes;   // the stash begins here

// grad[i] += (detw[i] + sum over CTAs of detb[cta][slot(i)]) * 2^-24 * scale: one thread per parameter, integer sums
__global__ void det_finalize_kernel(const long long* __restrict__ detw, const long long* __restrict__ detb, int nctas,
                                    float* __restrict__ grad, float out_scale, const float* __restrict__ sched,
                                    float inv_scale) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= PNTF_PARAM_COUNT) return;
  long long q = detw[i];
  int slot = -1;
  for (int l = 0; l <= 12; ++l)
    if (i >= c_b[l] && i < c_b[l] + c_M[l]) slot = c_db[l] + (i - c_b[l]);
  if (i >= c_w[13] && i < c_w[13] + 128) slot = kNumBiasSlots + (i - c_w[13]);
  if (i == c_b[13]) slot = kNumBiasSlots + 128;
  if (slot >= 0)
    for (int c = 0; c < nctas; ++c) q += detb[size_t(c) * kDetSlots + slot];
  if (q == 0) return;
  const float sc = sched != nullptr ? sched[1] * inv_scale : out_scale;
  grad[i] += float(double(q) * (1.0 / 16777216.0)) * sc;
}

size_t tc_workspace_bytes(int op, int E, int64_t N, int d, size_t budget_bytes) {
  if (op == PNTF_OP_PLAN) return align256(kFwdImgBytes) + 2048 + align256(size_t(N) * 4) + 256;   // + counters + hand-out order
  if (op != PNTF_OP_LOSS_BWD) return align256(kFwdImgBytes) + 512;
  const size_t P = tc_tile_pairs(d);
  const size_t ntiles = (size_t(E) * size_t(N) + P - 1) / P;
  if (budget_bytes == 0) budget_bytes = size_t(16) << 30;
  size_t tiles = budget_bytes / stash_tile_bytes();
  if (tiles > ntiles) tiles = ntiles;
  if (tiles < 1) tiles = 1;
  return kHeadBytes + tiles * stash_tile_bytes() + 256;
}

static void tc_carve_stash(TcStash& st, uint8_t* base, size_t tiles) {
  size_t off = 0;
  for (int l = 0; l <= 13; ++l) {
    if (stash_regen_layer(l)) { st.X[l] = nullptr; continue; }
    st.X[l] = base + off; off += tiles * layer_desc(l).K * 256;
  }
  for (int l = 0; l <= 12; ++l) { st.G[l] = base + off; off += tiles * layer_desc(l).M * 256; }
  st.G[13] = nullptr;
  st.Y5 = base + off; off += tiles * kHalfTileBytes;
  st.y13 = reinterpret_cast<float*>(base + off);
}
// per-(tile, layer) flags of the fused backward launch: behind the y13 rows
static uint32_t* stash_flags(const TcStash& st, size_t tiles) { return reinterpret_cast<uint32_t*>(st.y13 + tiles * kRows); }

// work units of the weight-gradient kernel: (layer, 128-feature half); every SM gets a CTA, shared out in proportion to
// the cost of a unit per tile: the KB it streams (32 KB of Ybar + K*256 B of X), or, for the units that regenerate their
// X operand (layers 0 and 6 with kStashRegen, tc_dw.cuh), a cost that stands for the generator's time per tile
// (PNTF_DW_GEN_COST, in the same KB units; measured default below)
static DwUnits make_dw_units(int sms) {
  static const int gen_cost = [] {
    const char* e = getenv("PNTF_DW_GEN_COST");
    const int v = e != nullptr ? atoi(e) : 0;
    return v > 0 ? v : 64;
  }();
  DwUnits u{};
  int n = 0;
  int cost[kDwMaxUnits], cnt[kDwMaxUnits];
  for (int l = 0; l <= 12; ++l)
    for (int h = 0; h < (layer_desc(l).M + 127) / 128; ++h) {
      u.l[n] = l; u.h[n] = h;
      cost[n] = stash_regen_layer(l) ? gen_cost : 32 + layer_desc(l).K / 4;
      cnt[n] = 1;
      ++n;
    }
  u.n = n;
  for (int extra = sms - n; extra > 0; --extra) {
    int best = 0;
    for (int i = 1; i < n; ++i)
      if (cost[i] * cnt[best] > cost[best] * cnt[i]) best = i;   // largest cost per CTA
    ++cnt[best];
  }
  int cta = 0;
  for (int i = 0; i < n; ++i) { u.cta_begin[i] = cta; cta += cnt[i]; }
  u.cta_begin[n] = cta;
  return u;
}

// Fused backward: `ncta` weight-gradient CTAs over the 19 (layer, half) units in proportion to the tensor time a unit
// needs per tile (K = 256: eight N = 256 UMMAs, K = 128: eight N = 128 UMMAs), at least one each.
static DwUnits make_dw_units_for(int ncta) {
  DwUnits u{};
  int n = 0;
  int cost[kDwMaxUnits], cnt[kDwMaxUnits];
  for (int l = 0; l <= 12; ++l)
    for (int h = 0; h < (layer_desc(l).M + 127) / 128; ++h) {
      u.l[n] = l; u.h[n] = h;
      cost[n] = layer_desc(l).K == 256 ? 2 : 1;
      cnt[n] = 1;
      ++n;
    }
  u.n = n;
  for (int extra = ncta - n; extra > 0; --extra) {
    int best = 0;
    for (int i = 1; i < n; ++i)
      if (cost[i] * cnt[best] > cost[best] * cnt[i]) best = i;   // largest cost per CTA
    ++cnt[best];
  }
  int cta = 0;
  for (int i = 0; i < n; ++i) { u.cta_begin[i] = cta; cta += cnt[i]; }
  u.cta_begin[n] = cta;
  return u;
}
// CTAs of the fused backward launch given to the weight-gradient role (0: separate launches, the default).
// PNTF_DW_CTAS=n turns the fused launch on with n weight-gradient CTAs.  Measured (profiles/r02a_*): correct (same parity
// tests) but slower than separate launches -- 66.9 ms against 30.0 + 16.8 ms per 1 Mi pairs with n = 48: a weight-gradient
// CTA holds only two 96 KB stages, runs latency-bound at ~2 900 clk per tile, falls behind the sweep and then reads from
// HBM after all.
static int fused_dw_ctas(int64_t tiles) {
  static const int forced = [] {
    const char* e = getenv("PNTF_DW_CTAS");
    return e != nullptr ? atoi(e) : 0;
  }();
  const int sms = sm_count();
  int n = forced;
  if (n <= 0) return 0;
  if (n < kDwMaxUnits) n = kDwMaxUnits;
  if (sms - n < 8 || tiles < 1) return 0;
  return n;
}

template <int D, int ORDER>
static int tc_loss_bwd(int variant, const float* params, const float* B, const float* coords, const float* speed, int E,
                       int64_t N, float alpha, float beta, float gamma, float* diff, float* grad_params, uint8_t* ws,
                       size_t ws_bytes, cudaStream_t s) {
  using C = TcCfg<D, ORDER>;
  const int64_t n_total = int64_t(E) * N;
  uint8_t* fimg = ws;
  uint8_t* rimg = ws + kFwdImgBytes;
  uint8_t* stash = ws + kHeadBytes;
  if (ws_bytes < kHeadBytes + stash_tile_bytes()) return fail(PNTF_ERR_WORKSPACE, "workspace too small for one tile");
  int64_t chunk_tiles = int64_t((ws_bytes - kHeadBytes) / stash_tile_bytes());
  const bool det = deterministic_mode() && sm_count() <= kDetMaxCtas;
  long long* detw = det ? reinterpret_cast<long long*>(ws + kImgBytes) : nullptr;
  long long* detb = det ? reinterpret_cast<long long*>(ws + kImgBytes + kDetWBytes) : nullptr;
  if (det) {
    cudaError_t e = cudaMemsetAsync(ws + kImgBytes, 0, kDetBytes, s);
    if (e != cudaSuccess) return fail(PNTF_ERR_CUDA, "cudaMemsetAsync(det): %s", cudaGetErrorString(e));
  }
  const int64_t ntiles_all = (n_total + C::P - 1) / C::P;
  if (chunk_tiles > ntiles_all) chunk_tiles = ntiles_all;
  else {
    // several chunks: equal sizes, rounded up to whole waves of the persistent kernels (every CTA of the forward and the
    // reverse sweep then owns the same number of tiles; only the last chunk is ragged)
    const int64_t sms = sm_count(), nchunks = (ntiles_all + chunk_tiles - 1) / chunk_tiles;
    int64_t per = (ntiles_all + nchunks - 1) / nchunks;
    per = (per + sms - 1) / sms * sms;
    if (per <= chunk_tiles) chunk_tiles = per;
    else if (chunk_tiles >= sms) chunk_tiles -= chunk_tiles % sms;
  }
  if (int rc = tc_pack(params, fimg, rimg, s)) return rc;
  auto kf = tc_forward_kernel<D, ORDER, true, TC_TAIL_LOSS>;
  auto kr = tc_reverse_kernel<D, ORDER>;
  auto kb = tc_backward_kernel<D, ORDER>;
  const size_t smem_f = kFwdSmem, smem_r = kRevSmem;
  const size_t smem_d = size_t(kDwStages) * kDwStageBytes + 1024;
  if (int rc = tc_set_smem(kf, smem_f)) return rc;
  if (int rc = tc_set_smem(kr, smem_r)) return rc;
  auto kd = tc_dw_kernel<D, ORDER>;
  if (int rc = tc_set_smem(kd, smem_d)) return rc;
  static_assert(kRevSmem >= size_t(kDwStages) * kDwStageBytes + 1024, "the fused launch uses the reverse sweep's shared memory size");
  if (int rc = tc_set_smem(kb, smem_r)) return rc;
  const int n_dw = fused_dw_ctas(chunk_tiles);
  const DwUnits units = n_dw > 0 ? make_dw_units_for(n_dw) : make_dw_units(sm_count());
  static uint32_t epoch = 0;
  const float out_scale = float(double(beta) / (double(E) * double(N)) / double(kGradScale));
  for (int64_t t0 = 0; t0 < ntiles_all; t0 += chunk_tiles) {
    const int64_t tiles = (ntiles_all - t0 < chunk_tiles) ? (ntiles_all - t0) : chunk_tiles;
    const int64_t begin = t0 * C::P;
    const int64_t count = (n_total - begin < tiles * C::P) ? (n_total - begin) : tiles * C::P;
    TcRevArgs r{};
    TcFwdArgs& a = r.f;
    a.params = params; a.wimg = fimg; a.B = B; a.coords = coords; a.speed = speed;
    apply_gather(a, D);
    apply_coords_bulk(a);
    a.n_total = n_total; a.n_per_env = N; a.pair_begin = begin; a.pair_count = count;
    a.diff = diff; a.single_variant = (variant == PNTF_LOSS_SINGLE); a.alpha = alpha; a.gamma = gamma;
    tc_carve_stash(a.st, stash, size_t(tiles));
    r.rimg = rimg; r.grad = grad_params; r.out_scale = out_scale;
    r.inv_scale = float(1.0 / (double(E) * double(N)) / double(kGradScale));
    r.detb = detb;
    const int grid = int(tiles < sm_count() ? tiles : sm_count());
    a.trace = (t0 == 0) ? trace_begin() : nullptr;
    prof_start(KC_TC_FORWARD, s);
    kf<<<grid, kTcThreads, smem_f, s>>>(a);
    prof_stop(s);
    trace_end(a.trace, s, "tc_forward(train)");
    a.trace = nullptr;
    if (int rc = launch_check("tc_forward(train)")) return rc;
    TcDwArgs dw{};
    dw.rg.B = a.B; dw.rg.coords = a.coords; dw.rg.index = a.index; dw.rg.cstride = a.cstride;
    dw.rg.n_total = a.n_total; dw.rg.n_per_env = a.n_per_env; dw.rg.pair_begin = a.pair_begin; dw.rg.pair_count = a.pair_count;
    dw.st = a.st; dw.units = units; dw.ntiles = tiles; dw.grad = grad_params; dw.out_scale = out_scale;
    dw.sched = a.sched; dw.inv_scale = r.inv_scale; dw.detw = detw;
    if (n_dw > 0) {
      // one launch: reverse sweep on the first CTAs, weight gradients trailing it on the last n_dw (tc_backward.cuh)
      uint32_t* flags = stash_flags(a.st, size_t(tiles));
      cudaError_t e = cudaMemsetAsync(flags, 0, size_t(tiles) * kFlagStride * sizeof(uint32_t), s);
      if (e != cudaSuccess) return fail(PNTF_ERR_CUDA, "cudaMemsetAsync(flags): %s", cudaGetErrorString(e));
      r.flags = flags; dw.flags = flags;
      r.epoch = dw.epoch = ++epoch == 0 ? ++epoch : epoch;
      const int n_rev = int(tiles < sm_count() - n_dw ? tiles : sm_count() - n_dw);
      a.trace = (t0 == 0) ? trace_begin() : nullptr;
      prof_start(KC_TC_BACKWARD, s);
      kb<<<n_rev + n_dw, kTcThreads, smem_r, s>>>(r, dw, n_rev);
      prof_stop(s);
      trace_end(a.trace, s, "tc_backward (role 1: acc half 0 ready, stash inputs ready, half 0 published, half 1 published)");
      a.trace = nullptr;
      if (int rc = launch_check("tc_backward")) return rc;
      continue;
    }
    a.trace = (t0 == 0) ? trace_begin() : nullptr;
    prof_start(KC_TC_REVERSE, s);
    kr<<<grid, kTcThreads, smem_r, s>>>(r);
    prof_stop(s);
    trace_end(a.trace, s, "tc_reverse (role 1: acc half 0 ready, stash inputs ready, half 0 published, half 1 published)");
    a.trace = nullptr;
    if (int rc = launch_check("tc_reverse")) return rc;
    prof_start(KC_TC_DW, s);
    kd<<<units.cta_begin[units.n], kDwThreads, smem_d, s>>>(dw);
    prof_stop(s);
    if (int rc = launch_check("tc_dw")) return rc;
  }
  if (det) {
    const float inv_scale = float(1.0 / (double(E) * double(N)) / double(kGradScale));
    det_finalize_kernel<<<(PNTF_PARAM_COUNT + 255) / 256, 256, 0, s>>>(detw, detb, sm_count(), grad_params, out_scale,
                                                                       g_sched, inv_scale);
    if (int rc = launch_check("det_finalize")) return rc;
  }
  return PNTF_OK;
}

// ---- debug: statistics of the fp16 adjoint tiles Ybar_l a pntf_loss_bwd call left in its workspace (tests only).
// Adjoints travel as fp16 scaled by kGradScale (tc_reverse.cuh); an overflow (inf) or a flush to zero would otherwise be
// silent.  out[0] = inf/NaN count, out[1] = subnormal count (0 < |x| < 2^-14), out[2] = nonzero count,
// out[3] = bits of the largest finite |x| (fp32).
__global__ void adjoint_stats_kernel(const __half* __restrict__ g, size_t n, unsigned long long* out) {
  unsigned long long bad = 0, sub = 0, nz = 0;
  float mx = 0.f;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x) {
    const float v = fabsf(__half2float(g[i]));
    if (!(v <= 65504.f)) { ++bad; continue; }
    if (v > 0.f) { ++nz; if (v < 6.103515625e-5f) ++sub; }
    mx = fmaxf(mx, v);
  }
  atomicAdd(out + 0, bad);
  atomicAdd(out + 1, sub);
  atomicAdd(out + 2, nz);
  atomicMax(reinterpret_cast<unsigned int*>(out + 3), __float_as_uint(mx));
}

// layer < 0: all adjoint tiles Ybar_0..12; layer = 100 + l: the forward operand tile X_l (l = 0..13) instead (X_0 and X_6
// are not stashed when kStashRegen: PNTF_ERR_ARG); layer = 200: the encoder outputs Y5, from which X_6 is regenerated
int tc_debug_adjoint_stats(const void* workspace, size_t workspace_bytes, int E, int64_t N, int d, int layer,
                           unsigned long long* out, void* stream) {
  const size_t P = tc_tile_pairs(d);
  const size_t ntiles = (size_t(E) * size_t(N) + P - 1) / P;
  if (workspace_bytes < kHeadBytes + ntiles * stash_tile_bytes())
    return fail(PNTF_ERR_WORKSPACE, "the adjoint statistics need a call that fitted one chunk (%zu tiles)", ntiles);
  TcStash st{};
  tc_carve_stash(st, const_cast<uint8_t*>(static_cast<const uint8_t*>(workspace)) + kHeadBytes, ntiles);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  cudaError_t e = cudaMemsetAsync(out, 0, 4 * sizeof(unsigned long long), s);
  if (e != cudaSuccess) return fail(PNTF_ERR_CUDA, "cudaMemsetAsync: %s", cudaGetErrorString(e));
  const uint8_t* base = st.G[0];
  size_t bytes = 0;
  if (layer == 200) {
    base = st.Y5;
    bytes = ntiles * size_t(kHalfTileBytes);
  } else if (layer >= 100 && layer <= 113) {
    if (stash_regen_layer(layer - 100))
      return fail(PNTF_ERR_ARG, "X_%d is not stashed: the weight-gradient kernel regenerates it", layer - 100);
    base = st.X[layer - 100];
    bytes = ntiles * size_t(layer_desc(layer - 100).K) * 256;
  } else if (layer >= 0 && layer <= 12) {
    base = st.G[layer];
    bytes = ntiles * size_t(layer_desc(layer).M) * 256;
  } else {
    for (int l = 0; l <= 12; ++l) bytes += ntiles * size_t(layer_desc(l).M) * 256;   // G[0] .. G[12] are contiguous
  }
  adjoint_stats_kernel<<<4 * sm_count(), 256, 0, s>>>(reinterpret_cast<const __half*>(base), bytes / 2, out);
  return launch_check("adjoint_stats");
}

static int tc_check_ws(void* workspace, size_t workspace_bytes, size_t need) {
  if (!workspace || workspace_bytes < need) return fail(PNTF_ERR_WORKSPACE, "workspace too small (%zu < %zu)", workspace_bytes, need);
  if ((reinterpret_cast<uintptr_t>(workspace) & 255) != 0) return fail(PNTF_ERR_ARG, "workspace must be 256-byte aligned");
  return PNTF_OK;
}

int tc_forward(int mode, int grad_flavour, const float* params, const float* B, const float* coords, int E, int64_t N,
               int d, float* tau, float* dtau, float* ltau, void* workspace, size_t workspace_bytes, void* stream) {
  if (N == 0) return PNTF_OK;
  if (int rc = tc_check_ws(workspace, workspace_bytes, kFwdImgBytes)) return rc;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  uint8_t* img = static_cast<uint8_t*>(workspace);
  if (int rc = tc_pack(params, img, nullptr, s)) return rc;
  TcFwdArgs a{};
  a.params = params; a.wimg = img; a.B = B; a.coords = coords;
  apply_gather(a, d);
  apply_coords_bulk(a);
  a.n_total = int64_t(E) * N; a.n_per_env = N; a.pair_begin = 0; a.pair_count = a.n_total;
  a.tau = tau; a.dtau = (mode >= PNTF_MODE_GRAD) ? dtau : nullptr; a.ltau = ltau;
  a.compat = (grad_flavour == PNTF_GRAD_COMPAT);
  // the value-only mode (NN.out) carries one tile row per endpoint: 64 pairs per 128-row tile (TcCfg ORDER 0)
  if (d == 3) {
    if (mode == PNTF_MODE_LAPLACE) return tc_launch_forward<3, 2, false, TC_TAIL_NONE>(a, s);
    if (mode == PNTF_MODE_VALUE) return tc_launch_forward<3, 0, false, TC_TAIL_NONE>(a, s);
    return tc_launch_forward<3, 1, false, TC_TAIL_NONE>(a, s);
  }
  if (mode == PNTF_MODE_LAPLACE) return tc_launch_forward<6, 2, false, TC_TAIL_NONE>(a, s);
  if (mode == PNTF_MODE_VALUE) return tc_launch_forward<6, 0, false, TC_TAIL_NONE>(a, s);
  return tc_launch_forward<6, 1, false, TC_TAIL_NONE>(a, s);
}

// Model.TravelTimes / Speed / Speed2 / Gradient on the tensor cores (forward kernel with the field tail)
int tc_field(int kind, int grad_flavour, const float* params, const float* B, const float* coords, int E, int64_t N,
             int d, float gamma, float* out, void* workspace, size_t workspace_bytes, void* stream) {
  if (N == 0) return PNTF_OK;
  if (int rc = tc_check_ws(workspace, workspace_bytes, kFwdImgBytes)) return rc;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  uint8_t* img = static_cast<uint8_t*>(workspace);
  if (int rc = tc_pack(params, img, nullptr, s)) return rc;
  TcFwdArgs a{};
  a.params = params; a.wimg = img; a.B = B; a.coords = coords;
  apply_gather(a, d);
  apply_coords_bulk(a);
  a.n_total = int64_t(E) * N; a.n_per_env = N; a.pair_begin = 0; a.pair_count = a.n_total;
  a.field = out; a.field_kind = kind; a.gamma = gamma;
  a.compat = (grad_flavour == PNTF_GRAD_COMPAT);
  // Speed2 uses the Laplacian only as the goal-side sum lap1 = ltau[:, dim:].sum(-1) (S:1228): summed second-order columns (TcCfg)
  const bool second = (kind == PNTF_FIELD_SPEED2);
  // TravelTimes needs tau only: value-only rows
  const bool value = (kind == PNTF_FIELD_TRAVEL_TIME);
  if (d == 3) {
    if (second) return tc_launch_forward<3, kOrderLapSum, false, TC_TAIL_FIELD>(a, s);
    if (value) return tc_launch_forward<3, 0, false, TC_TAIL_FIELD>(a, s);
    return tc_launch_forward<3, 1, false, TC_TAIL_FIELD>(a, s);
  }
  if (second) return tc_launch_forward<6, kOrderLapSum, false, TC_TAIL_FIELD>(a, s);
  if (value) return tc_launch_forward<6, 0, false, TC_TAIL_FIELD>(a, s);
  return tc_launch_forward<6, 1, false, TC_TAIL_FIELD>(a, s);
}

// ---- hand-out order of the path queries: longest start-goal distance first.  A query needs about |x1 - x0| / (2 eta)
// steps and a pass over a tile takes the same time however many of its slots are live, so the run ends with a tail of
// nearly empty passes as long as the longest query; starting the long queries first (longest-processing-time-first
// scheduling) leaves only short ones for the tail.  Counting sort by quantised distance, 256 buckets; the order inside
// a bucket is arbitrary, which cannot change results (rows of a tile are independent; tests/test_gpu_fullsize.py).
constexpr int kPlanBuckets = 256;
template <int D>
__device__ __forceinline__ int plan_bucket(const float* sg, int64_t q) {
  float d2 = 0.f;
#pragma unroll
  for (int i = 0; i < D; ++i) { const float v = __ldg(sg + q * 2 * D + D + i) - __ldg(sg + q * 2 * D + i); d2 = fmaf(v, v, d2); }
  const int b = int(sqrtf(d2) * (float(kPlanBuckets) / 2.5f));     // the unit box has diameter sqrt(6) < 2.5
  return kPlanBuckets - 1 - (b < 0 ? 0 : (b > kPlanBuckets - 1 ? kPlanBuckets - 1 : b));   // descending distance
}
template <int D>
__global__ void plan_hist_kernel(const float* __restrict__ sg, int64_t Q, int* __restrict__ hist) {
  const int64_t q = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (q < Q) atomicAdd(hist + plan_bucket<D>(sg, q), 1);
}
__global__ void plan_scan_kernel(int* hist) {      // exclusive scan of 256 counters in place (one block of 256 threads)
  __shared__ int s[kPlanBuckets];
  const int t = threadIdx.x;
  s[t] = hist[t];
  __syncthreads();
  for (int o = 1; o < kPlanBuckets; o <<= 1) {
    const int v = t >= o ? s[t - o] : 0;
    __syncthreads();
    s[t] += v;
    __syncthreads();
  }
  hist[t] = t ? s[t - 1] : 0;
}
template <int D>
__global__ void plan_scatter_kernel(const float* __restrict__ sg, int64_t Q, int* __restrict__ cursor, int32_t* __restrict__ order) {
  const int64_t q = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (q < Q) order[atomicAdd(cursor + plan_bucket<D>(sg, q), 1)] = int32_t(q);
}

// Batched path integration on the tensor cores: first-order forward per step, queries handed out per tile.
int tc_plan(const float* params, const float* B, const float* start_goal, int64_t Q, int d, float eta, float thr,
            int max_iter, int grad_flavour, int fill_tail, float* waypoints, int32_t* nsteps, void* workspace,
            size_t workspace_bytes, void* stream) {
  if (Q == 0) return PNTF_OK;
  if (int rc = tc_check_ws(workspace, workspace_bytes, align256(kFwdImgBytes) + 256)) return rc;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  uint8_t* img = static_cast<uint8_t*>(workspace);
  int* counter = reinterpret_cast<int*>(img + align256(kFwdImgBytes));        // [0]: query counter, [64 .. 320): buckets
  if (int rc = tc_pack(params, img, nullptr, s)) return rc;
  cudaError_t e = cudaMemsetAsync(counter, 0, 2048, s);
  if (e != cudaSuccess) return fail(PNTF_ERR_CUDA, "cudaMemsetAsync: %s", cudaGetErrorString(e));
  // hand-out order (needs 4 Q bytes of workspace behind the counters; dense queries only)
  int32_t* order = nullptr;
  const size_t order_off = align256(kFwdImgBytes) + 2048;
  static const bool order_off_env = [] { const char* e = getenv("PNTF_PLAN_ORDER"); return e != nullptr && atoi(e) == 0; }();
  if (!order_off_env && g_gather.index == nullptr && Q < (int64_t(1) << 31) && workspace_bytes >= order_off + size_t(Q) * 4) {
    order = reinterpret_cast<int32_t*>(img + order_off);
    int* hist = counter + 64;
    const unsigned nb = unsigned((Q + 255) / 256);
    if (d == 3) plan_hist_kernel<3><<<nb, 256, 0, s>>>(start_goal, Q, hist); else plan_hist_kernel<6><<<nb, 256, 0, s>>>(start_goal, Q, hist);
    plan_scan_kernel<<<1, kPlanBuckets, 0, s>>>(hist);
    if (d == 3) plan_scatter_kernel<3><<<nb, 256, 0, s>>>(start_goal, Q, hist, order); else plan_scatter_kernel<6><<<nb, 256, 0, s>>>(start_goal, Q, hist, order);
    if (int rc = launch_check("plan_order")) return rc;
  }
  TcFwdArgs a{};
  a.params = params; a.wimg = img; a.B = B; a.coords = start_goal;
  apply_gather(a, d);
  apply_coords_bulk(a);
  a.n_total = Q; a.n_per_env = Q; a.pair_begin = 0; a.pair_count = Q;
  a.compat = (grad_flavour == PNTF_GRAD_COMPAT);
  a.eta = eta; a.thr = thr; a.max_iter = max_iter; a.waypoints = waypoints; a.nsteps = nsteps; a.tile_counter = counter;
  a.fill_tail = fill_tail;
  a.order = order;
  const int64_t P = (d == 3) ? TcCfg<3, 1>::P : TcCfg<6, 1>::P;
  const int64_t ntiles = (Q + P - 1) / P;
  const int grid = int(ntiles < sm_count() ? ntiles : sm_count());
  prof_start(KC_PLAN, s);
  if (d == 3) {
    auto k = tc_forward_kernel<3, 1, false, TC_TAIL_PLAN>;
    if (int rc = tc_set_smem(k, kFwdSmem)) return rc;
    k<<<grid, kTcThreads, kFwdSmem, s>>>(a);
  } else {
    auto k = tc_forward_kernel<6, 1, false, TC_TAIL_PLAN>;
    if (int rc = tc_set_smem(k, kFwdSmem)) return rc;
    k<<<grid, kTcThreads, kFwdSmem, s>>>(a);
  }
  prof_stop(s);
  return launch_check("tc_plan");
}

int tc_loss(int with_grad, int variant, const float* params, const float* B, const float* coords, const float* speed,
            int E, int64_t N, int d, float alpha, float beta, float gamma, float* diff, float* loss_out,
            float* grad_params, void* workspace, size_t workspace_bytes, void* stream) {
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (with_grad) {
    if (int rc = tc_check_ws(workspace, workspace_bytes, kHeadBytes + stash_tile_bytes())) return rc;
#define PNTF_BWD(DD, OO) tc_loss_bwd<DD, OO>(variant, params, B, coords, speed, E, N, alpha, beta, gamma, diff, grad_params, \
                                            static_cast<uint8_t*>(workspace), workspace_bytes, s)
    int rc;
    if (loss_order() == kOrderLapSum) rc = (d == 3) ? PNTF_BWD(3, kOrderLapSum) : PNTF_BWD(6, kOrderLapSum);
    else rc = (d == 3) ? PNTF_BWD(3, 2) : PNTF_BWD(6, 2);
#undef PNTF_BWD
    if (rc) return rc;
    return finalize_loss(variant, diff, int64_t(E) * N, B, E, d, beta, loss_out, s);
  }
  if (int rc = tc_check_ws(workspace, workspace_bytes, kFwdImgBytes)) return rc;
  uint8_t* img = static_cast<uint8_t*>(workspace);
  if (int rc = tc_pack(params, img, nullptr, s)) return rc;
  TcFwdArgs a{};
  a.params = params; a.wimg = img; a.B = B; a.coords = coords; a.speed = speed;
  apply_gather(a, d);
  apply_coords_bulk(a);
  a.n_total = int64_t(E) * N; a.n_per_env = N; a.pair_begin = 0; a.pair_count = a.n_total;
  a.diff = diff; a.single_variant = (variant == PNTF_LOSS_SINGLE); a.alpha = alpha; a.gamma = gamma;
  int rc;
  if (loss_order() == kOrderLapSum)
    rc = (d == 3) ? tc_launch_forward<3, kOrderLapSum, false, TC_TAIL_LOSS>(a, s)
                  : tc_launch_forward<6, kOrderLapSum, false, TC_TAIL_LOSS>(a, s);
  else
    rc = (d == 3) ? tc_launch_forward<3, 2, false, TC_TAIL_LOSS>(a, s) : tc_launch_forward<6, 2, false, TC_TAIL_LOSS>(a, s);
  if (rc) return rc;
  return finalize_loss(variant, diff, int64_t(E) * N, B, E, d, beta, loss_out, s);
}

// ---- allocation-free path integration: scratch in a file-scope device array --------------------------------------------
// pntf_plan_static (include/pntf_b200.h) is pntf_plan(prec = PNTF_PREC_TC) for callers that cannot allocate (code under
// CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc).  The planner's scratch -- the fp16 weight
// image, the query counter + distance buckets, the hand-out order of the queries -- lives in g_plan_scratch instead of a
// caller-provided workspace.  It is sized for PNTF_PLAN_STATIC_MAX_QUERIES queries (1.31 MiB) and every GPU that loads the
// library holds it until the process exits.
//
// In host code the NAME g_plan_scratch is a host-side stand-in for the symbol, not a device address: a kernel handed
// `g_plan_scratch` from the host faults (or, where the GPU can reach host memory, silently uses the wrong memory).  The
// device address is asked for once per device with cudaGetSymbolAddress, with that device current, and only that pointer
// is passed on.
constexpr size_t kPlanStaticBytes =
    (kFwdImgBytes + 255) / 256 * 256 + 2048 + size_t(PNTF_PLAN_STATIC_MAX_QUERIES) * 4 + 256;
static __device__ __align__(256) unsigned char g_plan_scratch[kPlanStaticBytes];

constexpr int kMaxDevices = 64;
static void* g_plan_scratch_addr[kMaxDevices];     // device address of g_plan_scratch on device i (nullptr: not asked yet)
static std::mutex g_plan_scratch_mutex;

static int plan_scratch_ptr(void** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return fail(PNTF_ERR_CUDA, "cudaGetDevice: %s", cudaGetErrorString(e));
  if (dev < 0 || dev >= kMaxDevices) return fail(PNTF_ERR_UNSUPPORTED, "device ordinal %d beyond %d", dev, kMaxDevices);
  std::lock_guard<std::mutex> lock(g_plan_scratch_mutex);
  if (g_plan_scratch_addr[dev] == nullptr) {
    void* ptr = nullptr;
    e = cudaGetSymbolAddress(&ptr, g_plan_scratch);
    if (e != cudaSuccess || ptr == nullptr)
      return fail(PNTF_ERR_CUDA, "cudaGetSymbolAddress(g_plan_scratch) on device %d: %s", dev, cudaGetErrorString(e));
    g_plan_scratch_addr[dev] = ptr;
  }
  *out = g_plan_scratch_addr[dev];
  return PNTF_OK;
}

}  // namespace pntf

extern "C" int pntf_plan_static(const float* params, const float* B, const float* start_goal, int64_t Q, int d, float eta,
                                float thr, int max_iter, int grad_flavour, float* waypoints, int32_t* nsteps,
                                void* stream) {
  using namespace pntf;
  if (!params || !B || (!start_goal && Q > 0)) return fail(PNTF_ERR_ARG, "null pointer argument");
  if (Q < 0) return fail(PNTF_ERR_ARG, "bad shape Q=%lld", (long long)Q);
  if (d != 3 && d != 6) return fail(PNTF_ERR_UNSUPPORTED, "dimension d=%d not built (supported: 3, 6)", d);
  if (!waypoints || !nsteps || max_iter < 0) return fail(PNTF_ERR_ARG, "bad plan argument");
  if (Q > PNTF_PLAN_STATIC_MAX_QUERIES)
    return fail(PNTF_ERR_WORKSPACE,
                "pntf_plan_static: Q=%lld exceeds the built-in scratch (PNTF_PLAN_STATIC_MAX_QUERIES = %d queries per call); "
                "split the batch or use pntf_plan with a workspace",
                (long long)Q, PNTF_PLAN_STATIC_MAX_QUERIES);
  if (Q == 0) return PNTF_OK;
  void* scratch = nullptr;
  if (int rc = plan_scratch_ptr(&scratch)) return rc;
  // tc_plan rewrites the weight image and the hand-out order in full and clears the counters / buckets on `stream`
  // before the kernels that read them, so nothing of the previous call's contents is used.
  return tc_plan(params, B, start_goal, Q, d, eta, thr, max_iter, grad_flavour, 1, waypoints, nsteps, scratch,
                 kPlanStaticBytes, stream);
}
