// Tensor-core (tcgen05) precision mode entry points used by capi.cu.
#pragma once
#include <cstddef>
#include <cstdint>

#include <cuda_runtime.h>

namespace pntf {

// helpers defined in capi.cu
enum { KC_PACK = 0, KC_FORWARD, KC_REVERSE, KC_DW, KC_FINALIZE, KC_ADAMW, KC_PLAN, KC_TC_FORWARD, KC_TC_REVERSE,
       KC_TC_DW, KC_TC_BACKWARD, KC_COUNT };
int fail(int code, const char* fmt, ...);
void prof_start(int cls, cudaStream_t s);
void prof_stop(cudaStream_t s);
int launch_check(const char* what);
int sm_count();
bool deterministic_mode();   // pntf_set_deterministic / PNTF_DETERMINISTIC
int finalize_loss(int variant, const float* diff, int64_t n_total, const float* B, int E, int d, float beta,
                  float* loss_out, cudaStream_t s);   // beta is read from g_sched when that is set

// Batch source of the current call: dense arrays, or rows of a GPU-resident dataset picked by an index array
// (pntf_loss_*_gather).  Set by the entry point around the call; one host thread per GPU drives the library.
struct PairGather {
  const int32_t* index = nullptr;
  int row_stride = 0;
};
extern thread_local PairGather g_gather;
// device-resident {alpha, beta, gamma} of the current call (pntf_loss_bwd_gather_sched), else null
extern thread_local const float* g_sched;
template <class A>
inline void apply_gather(A& a, int d) {
  a.sched = g_sched;
  a.index = g_gather.index;
  a.cstride = g_gather.index ? g_gather.row_stride : 2 * d;
  a.sstride = g_gather.index ? g_gather.row_stride : 2;
}
// tensor-core forward: dense, 16-byte aligned coordinates are streamed into shared memory by bulk copies
template <class A>
inline void apply_coords_bulk(A& a) {
  a.coords_bulk = (a.index == nullptr && (reinterpret_cast<uintptr_t>(a.coords) & 15) == 0) ? 1 : 0;
}

size_t tc_workspace_bytes(int op, int E, int64_t N, int d, size_t budget_bytes);

int tc_forward(int mode, int grad_flavour, const float* params, const float* B, const float* coords, int E, int64_t N,
               int d, float* tau, float* dtau, float* ltau, void* workspace, size_t workspace_bytes, void* stream);

int tc_field(int kind, int grad_flavour, const float* params, const float* B, const float* coords, int E, int64_t N,
             int d, float gamma, float* out, void* workspace, size_t workspace_bytes, void* stream);

int tc_plan(const float* params, const float* B, const float* start_goal, int64_t Q, int d, float eta, float thr,
            int max_iter, int grad_flavour, int fill_tail, float* waypoints, int32_t* nsteps, void* workspace,
            size_t workspace_bytes, void* stream);

int tc_debug_adjoint_stats(const void* workspace, size_t workspace_bytes, int E, int64_t N, int d, int layer,
                           unsigned long long* out, void* stream);

// with_grad = 0: Model.Loss forward; 1: forward + parameter gradients
int tc_loss(int with_grad, int variant, const float* params, const float* B, const float* coords, const float* speed,
            int E, int64_t N, int d, float alpha, float beta, float gamma, float* diff, float* loss_out,
            float* grad_params, void* workspace, size_t workspace_bytes, void* stream);

}  // namespace pntf
