// Host side of the tensor-core precision mode (PNTF_PREC_TC): workspace carving and launches of the
// tcgen05 kernels.  Called from the C-ABI entry points in capi.cu.
#include "../../include/pntf_b200.h"
#include "tc_api.cuh"
#include "tc_forward.cuh"

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

constexpr size_t fwd_smem_bytes(bool train) { return size_t(train ? 2 : 1) * kTileBytes + size_t(train ? 5 : 8) * kStageBytes + 1024; }

template <int D, int ORDER, bool TRAIN, int TAIL>
static int tc_launch_forward(const TcFwdArgs& a, cudaStream_t s) {
  using C = TcCfg<D, ORDER>;
  auto k = tc_forward_kernel<D, ORDER, TRAIN, TAIL>;
  const size_t smem = fwd_smem_bytes(TRAIN);
  if (int rc = tc_set_smem(k, smem)) return rc;
  const int64_t ntiles = (a.pair_count + C::P - 1) / C::P;
  const int grid = int(ntiles < sm_count() ? ntiles : sm_count());
  prof_start(KC_TC_FORWARD, s);
  k<<<grid, kTcThreads, smem, s>>>(a);
  prof_stop(s);
  return launch_check("tc_forward");
}

size_t tc_workspace_bytes(int op, int E, int64_t N, int d, size_t budget_bytes) {
  (void)E; (void)N; (void)d; (void)budget_bytes;
  if (op == PNTF_OP_LOSS_BWD) return 0;   // reverse kernels: not built yet
  return align256(kFwdImgBytes) + 256;
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
  a.n_total = int64_t(E) * N; a.n_per_env = N; a.pair_begin = 0; a.pair_count = a.n_total;
  a.tau = tau; a.dtau = (mode >= PNTF_MODE_GRAD) ? dtau : nullptr; a.ltau = ltau;
  a.compat = (grad_flavour == PNTF_GRAD_COMPAT);
  // the value-only mode runs the first-order kernel (its extra columns are dropped)
  if (d == 3) {
    if (mode == PNTF_MODE_LAPLACE) return tc_launch_forward<3, 2, false, TC_TAIL_NONE>(a, s);
    return tc_launch_forward<3, 1, false, TC_TAIL_NONE>(a, s);
  }
  if (mode == PNTF_MODE_LAPLACE) return tc_launch_forward<6, 2, false, TC_TAIL_NONE>(a, s);
  return tc_launch_forward<6, 1, false, TC_TAIL_NONE>(a, s);
}

int tc_loss(int with_grad, int variant, const float* params, const float* B, const float* coords, const float* speed,
            int E, int64_t N, int d, float alpha, float beta, float gamma, float* diff, float* loss_out,
            float* grad_params, void* workspace, size_t workspace_bytes, void* stream) {
  (void)grad_params;
  if (with_grad) return fail(PNTF_ERR_UNSUPPORTED, "tensor-core reverse pass not built yet");
  if (int rc = tc_check_ws(workspace, workspace_bytes, kFwdImgBytes)) return rc;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  uint8_t* img = static_cast<uint8_t*>(workspace);
  if (int rc = tc_pack(params, img, nullptr, s)) return rc;
  TcFwdArgs a{};
  a.params = params; a.wimg = img; a.B = B; a.coords = coords; a.speed = speed;
  a.n_total = int64_t(E) * N; a.n_per_env = N; a.pair_begin = 0; a.pair_count = a.n_total;
  a.diff = diff; a.single_variant = (variant == PNTF_LOSS_SINGLE); a.alpha = alpha; a.gamma = gamma;
  int rc = (d == 3) ? tc_launch_forward<3, 2, false, TC_TAIL_LOSS>(a, s) : tc_launch_forward<6, 2, false, TC_TAIL_LOSS>(a, s);
  if (rc) return rc;
  return finalize_loss(variant, diff, int64_t(E) * N, B, E, d, beta, loss_out, s);
}

}  // namespace pntf
