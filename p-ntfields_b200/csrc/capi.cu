// C-ABI entry points of libpntf_b200.so (declared in include/pntf_b200.h): argument validation,
// workspace carving, kernel launches.  No torch types, no allocation, no host synchronisation.
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <atomic>

#include "../../include/pntf_b200.h"
#include "simt_misc.cuh"
#include "simt_reverse.cuh"
#include "tc_api.cuh"

namespace pntf {

static thread_local char g_err[512] = "";
thread_local PairGather g_gather;
thread_local const float* g_sched = nullptr;
static std::atomic<int64_t> g_launches{0};

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

#define PNTF_CUDA(expr)                                                                      \
  do {                                                                                       \
    cudaError_t _e = (expr);                                                                 \
    if (_e != cudaSuccess) return fail(PNTF_ERR_CUDA, "%s: %s", #expr, cudaGetErrorString(_e)); \
  } while (0)

// ---- optional per-kernel-class timing with CUDA events on the launching stream (bench.py) ----
struct ProfRec { int cls; cudaEvent_t a, b; };
constexpr int kProfMax = 8192;
static bool g_prof_on = false;
static int g_prof_n = 0;
static ProfRec g_prof[kProfMax];
static int g_prof_created = 0;

void prof_start(int cls, cudaStream_t s) {
  if (!g_prof_on || g_prof_n >= kProfMax) return;
  ProfRec& r = g_prof[g_prof_n];
  if (g_prof_n >= g_prof_created) {
    cudaEventCreate(&r.a);
    cudaEventCreate(&r.b);
    g_prof_created = g_prof_n + 1;
  }
  r.cls = cls;
  cudaEventRecord(r.a, s);
}
void prof_stop(cudaStream_t s) {
  if (!g_prof_on || g_prof_n >= kProfMax) return;
  cudaEventRecord(g_prof[g_prof_n].b, s);
  ++g_prof_n;
}

int launch_check(const char* what) {
  g_launches.fetch_add(1, std::memory_order_relaxed);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(PNTF_ERR_CUDA, "launch %s: %s", what, cudaGetErrorString(e));
  return PNTF_OK;
}

static int g_deterministic = -1;
bool deterministic_mode() {
  if (g_deterministic < 0) {
    const char* e = getenv("PNTF_DETERMINISTIC");
    g_deterministic = (e != nullptr && atoi(e) != 0) ? 1 : 0;
  }
  return g_deterministic == 1;
}

int sm_count() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 148;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
  }
  return n;
}

constexpr size_t kPackedBytes = ((size_t(kPackedCount) * 4 + 255) / 256) * 256;

static int tile_pairs(int d) { return d <= 3 ? 8 : 4; }

// floats of training stash per pair (X, Y, G for every layer), order-2 columns
static size_t stash_floats_per_pair(int d) {
  const size_t ce = 1 + 2 * d, cg = 1 + 4 * d;
  size_t s = 0;
  for (int l = 0; l < kNumLayers; ++l) {
    const LayerDesc L = layer_desc(l);
    const size_t rows = (l < 6) ? 2 * ce : cg;
    s += rows * (size_t(L.K) + 2 * size_t(L.M));
  }
  return s;
}

static void carve_stash(Stash& st, float* base, int d, int64_t chunk_pairs) {
  const size_t ce = 1 + 2 * d, cg = 1 + 4 * d;
  size_t off = 0;
  for (int l = 0; l < kNumLayers; ++l) {
    const LayerDesc L = layer_desc(l);
    const size_t rows = ((l < 6) ? 2 * ce : cg) * size_t(chunk_pairs);
    st.X[l] = base + off; off += rows * L.K;
    st.Y[l] = base + off; off += rows * L.M;
    st.G[l] = base + off; off += rows * L.M;
  }
}

template <typename K>
static int set_smem(K kernel, size_t bytes) {
  cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(bytes));
  if (e != cudaSuccess) return fail(PNTF_ERR_CUDA, "cudaFuncSetAttribute(smem=%zu): %s", bytes, cudaGetErrorString(e));
  return PNTF_OK;
}

static int pack(const float* params, float* fwd, float* bwd, int d, cudaStream_t s) {
  prof_start(KC_PACK, s);
  pack_weights_kernel<<<(kPackedCount + 255) / 256, 256, 0, s>>>(params, fwd, bwd, kThreads / tile_pairs(d));
  prof_stop(s);
  return launch_check("pack_weights");
}

template <int D, int ORDER, bool STASH, int TAIL>
static int launch_forward(const FwdArgs& a, cudaStream_t s) {
  using C = Cfg<D, ORDER>;
  auto k = simt_forward_kernel<D, ORDER, STASH, TAIL>;
  if (int rc = set_smem(k, C::SMEM_BYTES)) return rc;
  const int64_t ntiles = (a.pair_count + C::P - 1) / C::P;
  const int grid = int(ntiles < sm_count() ? ntiles : sm_count());
  prof_start(KC_FORWARD, s);
  k<<<grid, kThreads, C::SMEM_BYTES, s>>>(a);
  prof_stop(s);
  return launch_check("simt_forward");
}

template <int D>
static int forward_dispatch(int order, int tail, const FwdArgs& a, cudaStream_t s) {
  if (tail == TAIL_LOSS) return launch_forward<D, 2, false, TAIL_LOSS>(a, s);
  if (tail == TAIL_FIELD) {
    if (order == 0) return launch_forward<D, 0, false, TAIL_FIELD>(a, s);
    if (order == 1) return launch_forward<D, 1, false, TAIL_FIELD>(a, s);
    return launch_forward<D, 2, false, TAIL_FIELD>(a, s);
  }
  if (order == 0) return launch_forward<D, 0, false, TAIL_NONE>(a, s);
  if (order == 1) return launch_forward<D, 1, false, TAIL_NONE>(a, s);
  return launch_forward<D, 2, false, TAIL_NONE>(a, s);
}

static int check_prec(int prec) {
  if (prec != PNTF_PREC_FP32 && prec != PNTF_PREC_TC) return fail(PNTF_ERR_ARG, "bad precision mode %d", prec);
  return PNTF_OK;
}
static int check_common(const void* params, const void* B, const void* coords, int E, int64_t N, int d) {
  if (!params || !B || (!coords && N > 0)) return fail(PNTF_ERR_ARG, "null pointer argument");
  if (E < 1 || N < 0) return fail(PNTF_ERR_ARG, "bad shape E=%d N=%lld", E, (long long)N);
  if (d != 3 && d != 6) return fail(PNTF_ERR_UNSUPPORTED, "dimension d=%d not built (supported: 3, 6)", d);
  return PNTF_OK;
}

template <int D>
static int loss_bwd_simt(int variant, const float* params, const float* B, const float* coords, const float* speed,
                         int E, int64_t N, float alpha, float beta, float gamma, float* diff, float* grad_params,
                         float* ws, size_t ws_bytes, cudaStream_t s) {
  using C = Cfg<D, 2>;
  const int64_t n_total = int64_t(E) * N;
  float* packed_f = ws;
  float* packed_b = ws + kPackedBytes / 4;
  float* stash_base = ws + 2 * (kPackedBytes / 4);
  const size_t avail = ws_bytes - 2 * kPackedBytes;
  const size_t per_pair = stash_floats_per_pair(D) * 4;
  int64_t chunk = int64_t(avail / per_pair) / C::P * C::P;
  const int64_t n_pad = (n_total + C::P - 1) / C::P * C::P;
  if (chunk > n_pad) chunk = n_pad;
  if (chunk < C::P) return fail(PNTF_ERR_WORKSPACE, "workspace too small for one tile");
  if (int rc = pack(params, packed_f, packed_b, D, s)) return rc;
  auto kf = simt_forward_kernel<D, 2, true, TAIL_NONE>;
  auto kr = simt_reverse_kernel<D>;
  if (int rc = set_smem(kf, C::SMEM_BYTES)) return rc;
  if (int rc = set_smem(kr, C::SMEM_BYTES)) return rc;
  for (int64_t begin = 0; begin < n_total; begin += chunk) {
    const int64_t count = (n_total - begin < chunk) ? (n_total - begin) : chunk;
    const int64_t count_pad = (count + C::P - 1) / C::P * C::P;
    BwdArgs b{};
    FwdArgs& a = b.f;
    a.params = params; a.packed = packed_f; a.B = B; a.coords = coords; a.speed = speed;
    apply_gather(a, D);
    a.n_total = n_total; a.n_per_env = N; a.pair_begin = begin; a.pair_count = count;
    a.diff = diff; a.single_variant = (variant == PNTF_LOSS_SINGLE); a.alpha = alpha; a.gamma = gamma;
    carve_stash(a.st, stash_base, D, count_pad);
    b.packed_bwd = packed_b;
    b.wscale = beta / float(double(E) * double(N));
    b.inv_count = float(1.0 / (double(E) * double(N)));
    const int64_t ntiles = count_pad / C::P;
    const int grid = int(ntiles < sm_count() ? ntiles : sm_count());
    prof_start(KC_FORWARD, s);
    kf<<<grid, kThreads, C::SMEM_BYTES, s>>>(a);
    prof_stop(s);
    if (int rc = launch_check("simt_forward(stash)")) return rc;
    prof_start(KC_REVERSE, s);
    kr<<<grid, kThreads, C::SMEM_BYTES, s>>>(b);
    prof_stop(s);
    if (int rc = launch_check("simt_reverse")) return rc;
    prof_start(KC_DW, s);
    for (int l = 0; l < kNumLayers; ++l) {
      const LayerDesc L = layer_desc(l);
      DwArgs dw{};
      dw.G = a.st.G[l]; dw.X = a.st.X[l];
      dw.dW = grad_params + L.w; dw.db = grad_params + L.b;
      dw.M = L.M; dw.K = L.K;
      dw.cols_per_point = (l < 6) ? C::CE : C::CG;
      dw.rows = ((l < 6) ? 2 * C::CE : C::CG) * count_pad;
      const int tm = (L.M + 127) / 128, tk = L.K / 128;
      int64_t splits = (2 * sm_count()) / (tm * tk);
      if (splits < 1) splits = 1;
      int64_t rps = (dw.rows + splits - 1) / splits;
      rps = (rps + 15) / 16 * 16;
      if (rps < 16) rps = 16;
      splits = (dw.rows + rps - 1) / rps;
      dw.rows_per_split = rps;
      dw_kernel<<<dim3(tm, tk, unsigned(splits)), 256, 0, s>>>(dw);
      if (int rc = launch_check("dw_kernel")) return rc;
    }
    prof_stop(s);
  }
  return PNTF_OK;
}

}  // namespace pntf

using namespace pntf;

extern "C" {

int pntf_version(void) { return 100; }
const char* pntf_last_error(void) { return g_err; }
int64_t pntf_launch_count(void) { return g_launches.load(); }

int pntf_set_deterministic(int on) {
  g_deterministic = on ? 1 : 0;
  return PNTF_OK;
}

int pntf_profile_begin(void) {
  g_prof_n = 0;
  g_prof_on = true;
  return PNTF_OK;
}

int pntf_profile_end(double* ms_by_class, int64_t* count_by_class, int n_classes) {
  g_prof_on = false;
  for (int i = 0; i < n_classes; ++i) { ms_by_class[i] = 0.0; count_by_class[i] = 0; }
  for (int i = 0; i < g_prof_n; ++i) {
    PNTF_CUDA(cudaEventSynchronize(g_prof[i].b));
    float ms = 0.f;
    PNTF_CUDA(cudaEventElapsedTime(&ms, g_prof[i].a, g_prof[i].b));
    if (g_prof[i].cls < n_classes) { ms_by_class[g_prof[i].cls] += ms; count_by_class[g_prof[i].cls] += 1; }
  }
  g_prof_n = 0;
  return PNTF_OK;
}

int pntf_param_offset(int index, int64_t* offset, int64_t* rows, int64_t* cols, int* used_by_forward) {
  if (index < 0 || index >= kNumTensors) return fail(PNTF_ERR_ARG, "tensor index %d out of range", index);
  const TensorDesc t = tensor_desc(index);
  if (offset) *offset = t.offset;
  if (rows) *rows = t.rows;
  if (cols) *cols = t.cols;
  if (used_by_forward) *used_by_forward = t.used;
  return PNTF_OK;
}

size_t pntf_workspace_bytes(int op, int E, int64_t N, int d, int prec, size_t budget_bytes) {
  if (d != 3 && d != 6) return 0;
  if (prec == PNTF_PREC_TC) return tc_workspace_bytes(op, E, N, d, budget_bytes);
  if (op != PNTF_OP_LOSS_BWD) return kPackedBytes + 256;
  const size_t P = tile_pairs(d);
  const size_t n_pad = (size_t(E) * size_t(N) + P - 1) / P * P;
  const size_t per_pair = stash_floats_per_pair(d) * 4;
  if (budget_bytes == 0) budget_bytes = size_t(8) << 30;
  size_t stash = n_pad * per_pair;
  if (stash > budget_bytes) stash = budget_bytes / (per_pair * P) * (per_pair * P);
  if (stash < per_pair * P) stash = per_pair * P;
  return 2 * kPackedBytes + stash + 256;
}

static int forward_common(int order, int tail, int field_kind, int grad_flavour, const float* params, const float* B,
                          const float* coords, const float* speed, int E, int64_t N, int d, float alpha, float gamma,
                          int single_variant, float* tau, float* dtau, float* ltau, float* field, float* diff,
                          void* workspace, size_t workspace_bytes, void* stream) {
  if (int rc = check_common(params, B, coords, E, N, d)) return rc;
  if (N == 0) return PNTF_OK;
  if (!workspace || workspace_bytes < kPackedBytes) return fail(PNTF_ERR_WORKSPACE, "workspace too small (%zu < %zu)", workspace_bytes, kPackedBytes);
  if ((reinterpret_cast<uintptr_t>(workspace) & 15) != 0) return fail(PNTF_ERR_ARG, "workspace must be 16-byte aligned");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  float* packed = static_cast<float*>(workspace);
  if (int rc = pack(params, packed, nullptr, d, s)) return rc;
  FwdArgs a{};
  a.params = params; a.packed = packed; a.B = B; a.coords = coords; a.speed = speed;
  apply_gather(a, d);
  a.n_total = int64_t(E) * N; a.n_per_env = N; a.pair_begin = 0; a.pair_count = a.n_total;
  a.tau = tau; a.dtau = dtau; a.ltau = ltau; a.field = field; a.field_kind = field_kind;
  a.diff = diff; a.single_variant = single_variant; a.alpha = alpha; a.gamma = gamma;
  a.compat = (grad_flavour == PNTF_GRAD_COMPAT);
  if (d == 3) return forward_dispatch<3>(order, tail, a, s);
  return forward_dispatch<6>(order, tail, a, s);
}

int pntf_forward(int mode, int grad_flavour, const float* params, const float* B, const float* coords, int E, int64_t N,
                 int d, float* tau, float* dtau, float* ltau, int prec, void* workspace, size_t workspace_bytes,
                 void* stream) {
  if (mode < 0 || mode > 2) return fail(PNTF_ERR_ARG, "bad mode %d", mode);
  if (int rc = check_prec(prec)) return rc;
  if (grad_flavour == PNTF_GRAD_COMPAT && mode != PNTF_MODE_GRAD)
    return fail(PNTF_ERR_ARG, "the compat gradient flavour exists for MODE_GRAD only");
  if (prec == PNTF_PREC_TC) {
    if (int rc = check_common(params, B, coords, E, N, d)) return rc;
    return tc_forward(mode, grad_flavour, params, B, coords, E, N, d, tau, dtau, ltau, workspace, workspace_bytes, stream);
  }
  return forward_common(mode, TAIL_NONE, 0, grad_flavour, params, B, coords, nullptr, E, N, d, 1.f, 0.f, 0, tau, dtau,
                        ltau, nullptr, nullptr, workspace, workspace_bytes, stream);
}

int pntf_field(int kind, int grad_flavour, const float* params, const float* B, const float* coords, int E, int64_t N,
               int d, float gamma, float* out, int prec, void* workspace, size_t workspace_bytes, void* stream) {
  if (kind < 0 || kind > 3 || !out) return fail(PNTF_ERR_ARG, "bad field kind %d or null output", kind);
  if (int rc = check_prec(prec)) return rc;
  if (grad_flavour == PNTF_GRAD_COMPAT && kind != PNTF_FIELD_DIRECTION)
    return fail(PNTF_ERR_ARG, "the compat gradient flavour exists for FIELD_DIRECTION only");
  const int order = (kind == PNTF_FIELD_TRAVEL_TIME) ? 0 : (kind == PNTF_FIELD_SPEED2 ? 2 : 1);
  if (prec == PNTF_PREC_TC) {
    if (int rc = check_common(params, B, coords, E, N, d)) return rc;
    return tc_field(kind, grad_flavour, params, B, coords, E, N, d, gamma, out, workspace, workspace_bytes, stream);
  }
  return forward_common(order, TAIL_FIELD, kind, grad_flavour, params, B, coords, nullptr, E, N, d, 1.f, gamma, 0,
                        nullptr, nullptr, nullptr, out, nullptr, workspace, workspace_bytes, stream);
}

}  // extern "C"
namespace pntf {
int finalize_loss(int variant, const float* diff, int64_t n_total, const float* B, int E, int d, float beta,
                  float* loss_out, cudaStream_t s) {
  prof_start(KC_FINALIZE, s);
  loss_finalize_kernel<<<1, 1024, 0, s>>>(diff, n_total, B, E * d * kNumFreq, variant == PNTF_LOSS_MULTI, beta, g_sched,
                                          loss_out);
  prof_stop(s);
  return launch_check("loss_finalize");
}
}  // namespace pntf
extern "C" {

int pntf_loss_fwd(int variant, const float* params, const float* B, const float* coords, const float* speed, int E,
                  int64_t N, int d, float alpha, float beta, float gamma, float* diff, float* loss_out, int prec,
                  void* workspace, size_t workspace_bytes, void* stream) {
  if (variant != PNTF_LOSS_SINGLE && variant != PNTF_LOSS_MULTI) return fail(PNTF_ERR_ARG, "bad loss variant");
  if (!speed || !diff || !loss_out) return fail(PNTF_ERR_ARG, "null speed/diff/loss_out");
  if (int rc = check_prec(prec)) return rc;
  if (int rc = check_common(params, B, coords, E, N, d)) return rc;
  if (N == 0) return fail(PNTF_ERR_ARG, "empty batch has no loss");
  if (prec == PNTF_PREC_TC)
    return tc_loss(0, variant, params, B, coords, speed, E, N, d, alpha, beta, gamma, diff, loss_out, nullptr,
                   workspace, workspace_bytes, stream);
  if (int rc = forward_common(2, TAIL_LOSS, 0, 0, params, B, coords, speed, E, N, d, alpha, gamma,
                              variant == PNTF_LOSS_SINGLE, nullptr, nullptr, nullptr, nullptr, diff, workspace,
                              workspace_bytes, stream))
    return rc;
  return finalize_loss(variant, diff, int64_t(E) * N, B, E, d, beta, loss_out, static_cast<cudaStream_t>(stream));
}

int pntf_loss_bwd(int variant, const float* params, const float* B, const float* coords, const float* speed, int E,
                  int64_t N, int d, float alpha, float beta, float gamma, float* diff, float* loss_out,
                  float* grad_params, int prec, void* workspace, size_t workspace_bytes, void* stream) {
  if (variant != PNTF_LOSS_SINGLE && variant != PNTF_LOSS_MULTI) return fail(PNTF_ERR_ARG, "bad loss variant");
  if (!speed || !diff || !loss_out || !grad_params) return fail(PNTF_ERR_ARG, "null speed/diff/loss_out/grad_params");
  if (int rc = check_prec(prec)) return rc;
  if (int rc = check_common(params, B, coords, E, N, d)) return rc;
  if (N == 0) return fail(PNTF_ERR_ARG, "empty batch has no loss");
  if (prec == PNTF_PREC_TC)
    return tc_loss(1, variant, params, B, coords, speed, E, N, d, alpha, beta, gamma, diff, loss_out, grad_params,
                   workspace, workspace_bytes, stream);
  if (!workspace || workspace_bytes < 2 * kPackedBytes + 4096) return fail(PNTF_ERR_WORKSPACE, "workspace too small");
  if ((reinterpret_cast<uintptr_t>(workspace) & 15) != 0) return fail(PNTF_ERR_ARG, "workspace must be 16-byte aligned");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  int rc;
  if (d == 3)
    rc = loss_bwd_simt<3>(variant, params, B, coords, speed, E, N, alpha, beta, gamma, diff, grad_params,
                          static_cast<float*>(workspace), workspace_bytes, s);
  else
    rc = loss_bwd_simt<6>(variant, params, B, coords, speed, E, N, alpha, beta, gamma, diff, grad_params,
                          static_cast<float*>(workspace), workspace_bytes, s);
  if (rc) return rc;
  return finalize_loss(variant, diff, int64_t(E) * N, B, E, d, beta, loss_out, s);
}

// ---- fused batch gather: the batch is index[E*N] rows of a GPU-resident dataset
struct GatherScope {
  GatherScope(const int32_t* index, int stride) { g_gather.index = index; g_gather.row_stride = stride; }
  ~GatherScope() { g_gather = PairGather{}; }
};
static int check_gather(const float* dataset, int64_t dataset_rows, const int32_t* index, int d) {
  if (!dataset || !index) return fail(PNTF_ERR_ARG, "null dataset/index");
  if (d != 3 && d != 6) return fail(PNTF_ERR_ARG, "d must be 3 or 6");
  if (dataset_rows <= 0 || dataset_rows > int64_t(0x7fffffff)) return fail(PNTF_ERR_ARG, "dataset_rows out of range");
  return PNTF_OK;
}

int pntf_loss_fwd_gather(int variant, const float* params, const float* B, const float* dataset, int64_t dataset_rows,
                         const int32_t* index, int E, int64_t N, int d, float alpha, float beta, float gamma, float* diff,
                         float* loss_out, int prec, void* workspace, size_t workspace_bytes, void* stream) {
  if (int rc = check_gather(dataset, dataset_rows, index, d)) return rc;
  GatherScope scope(index, 2 * d + 2);
  return pntf_loss_fwd(variant, params, B, dataset, dataset + 2 * d, E, N, d, alpha, beta, gamma, diff, loss_out, prec,
                       workspace, workspace_bytes, stream);
}

int pntf_loss_bwd_gather(int variant, const float* params, const float* B, const float* dataset, int64_t dataset_rows,
                         const int32_t* index, int E, int64_t N, int d, float alpha, float beta, float gamma, float* diff,
                         float* loss_out, float* grad_params, int prec, void* workspace, size_t workspace_bytes,
                         void* stream) {
  if (int rc = check_gather(dataset, dataset_rows, index, d)) return rc;
  GatherScope scope(index, 2 * d + 2);
  return pntf_loss_bwd(variant, params, B, dataset, dataset + 2 * d, E, N, d, alpha, beta, gamma, diff, loss_out,
                       grad_params, prec, workspace, workspace_bytes, stream);
}

// ---- device-resident schedule: an epoch of training steps can be captured in a CUDA graph and replayed
struct SchedScope {
  explicit SchedScope(const float* s) { g_sched = s; }
  ~SchedScope() { g_sched = nullptr; }
};

int pntf_loss_bwd_gather_sched(int variant, const float* params, const float* B, const float* dataset, int64_t dataset_rows,
                               const int32_t* index, int E, int64_t N, int d, const float* sched, float* diff,
                               float* loss_out, float* grad_params, int prec, void* workspace, size_t workspace_bytes,
                               void* stream) {
  if (!sched) return fail(PNTF_ERR_ARG, "null schedule");
  SchedScope scope(sched);
  return pntf_loss_bwd_gather(variant, params, B, dataset, dataset_rows, index, E, N, d, 1.f, 1.f, 0.f, diff, loss_out,
                              grad_params, prec, workspace, workspace_bytes, stream);
}

int pntf_adamw_sched(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, const uint8_t* mask, int64_t n,
                     const float* lr_step, double beta1, double beta2, double eps, double weight_decay, double grad_scale,
                     void* stream) {
  if (!params || !grads || !exp_avg || !exp_avg_sq || !lr_step || n < 0) return fail(PNTF_ERR_ARG, "bad adamw argument");
  if (n == 0) return PNTF_OK;
  AdamArgs a{};
  a.p = params; a.g = grads; a.m = exp_avg; a.v = exp_avg_sq; a.mask = mask; a.n = n;
  a.omb1 = float(1.0 - beta1); a.b2 = float(beta2); a.omb2 = float(1.0 - beta2); a.eps = float(eps);
  a.grad_scale = float(grad_scale);
  a.lr_step = lr_step; a.b1d = beta1; a.b2d = beta2; a.wdd = weight_decay;
  const int blocks = int((n + 255) / 256 < 4 * sm_count() ? (n + 255) / 256 : 4 * sm_count());
  prof_start(KC_ADAMW, static_cast<cudaStream_t>(stream));
  adamw_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(a);
  prof_stop(static_cast<cudaStream_t>(stream));
  return launch_check("adamw(sched)");
}

int pntf_debug_adjoint_stats(const void* workspace, size_t workspace_bytes, int E, int64_t N, int d, int layer,
                             uint64_t* out, void* stream) {
  if (!workspace || !out || E < 1 || N < 1 || (d != 3 && d != 6)) return fail(PNTF_ERR_ARG, "bad adjoint-statistics argument");
  return tc_debug_adjoint_stats(workspace, workspace_bytes, E, N, d, layer, reinterpret_cast<unsigned long long*>(out), stream);
}

int pntf_adamw(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, const uint8_t* mask, int64_t n,
               double lr, double beta1, double beta2, double eps, double weight_decay, int64_t step,
               double grad_scale, void* stream) {
  if (!params || !grads || !exp_avg || !exp_avg_sq || n < 0 || step < 1) return fail(PNTF_ERR_ARG, "bad adamw argument");
  if (n == 0) return PNTF_OK;
  AdamArgs a{};
  a.p = params; a.g = grads; a.m = exp_avg; a.v = exp_avg_sq; a.mask = mask; a.n = n;
  a.decay = float(1.0 - lr * weight_decay);
  a.omb1 = float(1.0 - beta1); a.b2 = float(beta2); a.omb2 = float(1.0 - beta2); a.eps = float(eps);
  a.step_size = float(lr / (1.0 - std::pow(beta1, double(step))));
  a.bc2_sqrt = float(std::sqrt(1.0 - std::pow(beta2, double(step))));
  a.grad_scale = float(grad_scale);
  const int blocks = int((n + 255) / 256 < 4 * sm_count() ? (n + 255) / 256 : 4 * sm_count());
  prof_start(KC_ADAMW, static_cast<cudaStream_t>(stream));
  adamw_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(a);
  prof_stop(static_cast<cudaStream_t>(stream));
  return launch_check("adamw");
}

static int plan_impl(const float* params, const float* B, const float* start_goal, int64_t Q, int d, float eta, float thr,
                     int max_iter, int grad_flavour, int fill_tail, float* waypoints, int32_t* nsteps, int prec,
                     void* workspace, size_t workspace_bytes, void* stream) {
  if (int rc = check_common(params, B, start_goal, 1, Q, d)) return rc;
  if (!waypoints || !nsteps || max_iter < 0) return fail(PNTF_ERR_ARG, "bad plan argument");
  if (int rc = check_prec(prec)) return rc;
  if (Q == 0) return PNTF_OK;
  if (prec == PNTF_PREC_TC)
    return tc_plan(params, B, start_goal, Q, d, eta, thr, max_iter, grad_flavour, fill_tail, waypoints, nsteps, workspace,
                   workspace_bytes, stream);
  if (!workspace || workspace_bytes < kPackedBytes) return fail(PNTF_ERR_WORKSPACE, "workspace too small");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  float* packed = static_cast<float*>(workspace);
  if (int rc = pack(params, packed, nullptr, d, s)) return rc;
  PlanArgs pa{};
  FwdArgs& a = pa.f;
  a.params = params; a.packed = packed; a.B = B; a.coords = start_goal;
  apply_gather(a, d);
  a.n_total = Q; a.n_per_env = Q; a.pair_begin = 0; a.pair_count = Q;
  a.compat = (grad_flavour == PNTF_GRAD_COMPAT);
  pa.eta = eta; pa.thr = thr; pa.max_iter = max_iter; pa.waypoints = waypoints; pa.nsteps = nsteps;
  pa.fill_tail = fill_tail;
  const int P = tile_pairs(d);
  const int64_t ntiles = (Q + P - 1) / P;
  // more CTAs than SMs do not help (1 CTA/SM); tiles are handed out round-robin
  const int grid = int(ntiles < sm_count() ? ntiles : sm_count());
  prof_start(KC_PLAN, s);
  if (d == 3) {
    auto k = simt_plan_kernel<3>;
    if (int rc = set_smem(k, Cfg<3, 1>::SMEM_BYTES)) return rc;
    k<<<grid, kThreads, Cfg<3, 1>::SMEM_BYTES, s>>>(pa);
  } else {
    auto k = simt_plan_kernel<6>;
    if (int rc = set_smem(k, Cfg<6, 1>::SMEM_BYTES)) return rc;
    k<<<grid, kThreads, Cfg<6, 1>::SMEM_BYTES, s>>>(pa);
  }
  prof_stop(s);
  return launch_check("simt_plan");
}

int pntf_plan(const float* params, const float* B, const float* start_goal, int64_t Q, int d, float eta, float thr,
              int max_iter, int grad_flavour, float* waypoints, int32_t* nsteps, int prec, void* workspace,
              size_t workspace_bytes, void* stream) {
  return plan_impl(params, B, start_goal, Q, d, eta, thr, max_iter, grad_flavour, 1, waypoints, nsteps, prec, workspace,
                   workspace_bytes, stream);
}

int pntf_plan_compact(const float* params, const float* B, const float* start_goal, int64_t Q, int d, float eta, float thr,
                      int max_iter, int grad_flavour, float* waypoints, int32_t* nsteps, int prec, void* workspace,
                      size_t workspace_bytes, void* stream) {
  return plan_impl(params, B, start_goal, Q, d, eta, thr, max_iter, grad_flavour, 0, waypoints, nsteps, prec, workspace,
                   workspace_bytes, stream);
}

}  // extern "C"
