/*
 * pntf_b200.h -- C-ABI of the B200-native P-NTFields hot path (libpntf_b200.so).
 *
 * The reference (yhsong0804/P-NTFields) is pure Python/PyTorch and has no FFI; its boundary for this
 * path is the Python class API of models/model_res_sigmoid.py (S:) and
 * models/model_res_sigmoid_multi.py (M:).  Each entry point below states which reference method(s)
 * it replaces.  The Python mirror of those classes lives in p-ntfields_b200/pntf and calls these
 * symbols through ctypes.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer into caller-owned memory (fp32, row-major, contiguous)
 *     unless its name starts with h_; the library never allocates, frees or keeps pointers;
 *   - `stream` is a cudaStream_t passed as void*; all work is enqueued on it, no host sync
 *     (except pntf_plan when h_sync != 0);
 *   - return value 0 = ok, nonzero = error (PNTF_ERR_*); pntf_last_error() returns a
 *     thread-local message;
 *   - d = C-space dimension (3 or 6), a start/goal pair is 2d floats (x0 then x1);
 *   - params = the flat fp32 parameter buffer in state_dict order of S:159-181 / M:153-175
 *     (576129 floats; see pntf_param_offset);
 *   - B = Fourier matrix [E, d, 128] (one per environment; single-env callers pass E = 1 and
 *     NN.B which is already [d,128], S:139).  Pair n of coords[E*N, 2d] uses B[n / N];
 *   - prec: PNTF_PREC_FP32 = fp32 SIMT verify mode, PNTF_PREC_TC = tcgen05 tensor-core mode.
 */
#ifndef PNTF_B200_H_
#define PNTF_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PNTF_PARAM_COUNT 576129
#define PNTF_NUM_TENSORS 30

enum { PNTF_OK = 0, PNTF_ERR_ARG = 1, PNTF_ERR_UNSUPPORTED = 2, PNTF_ERR_WORKSPACE = 3, PNTF_ERR_CUDA = 4 };
enum { PNTF_PREC_FP32 = 0, PNTF_PREC_TC = 1 };
/* derivative order of the forward pass: NN.out / NN.out_grad / NN.out_laplace */
enum { PNTF_MODE_VALUE = 0, PNTF_MODE_GRAD = 1, PNTF_MODE_LAPLACE = 2 };
/* Model.Loss flavour: S:869-936 (square roots on the ratio) / M:897-951 (none, + 0.01|B|^2) */
enum { PNTF_LOSS_SINGLE = 0, PNTF_LOSS_MULTI = 1 };
/* pntf_field kinds */
enum {
  PNTF_FIELD_TRAVEL_TIME = 0, /* Model.TravelTimes  S:1169-1183, M:1173-1186 */
  PNTF_FIELD_SPEED = 1,       /* Model.Speed        S:1192-1216, M:1195-1216 */
  PNTF_FIELD_SPEED2 = 2,      /* Model.Speed2       S:1218-1245 (uses gamma)  */
  PNTF_FIELD_DIRECTION = 3    /* Model.Gradient     S:1247-1282, M:1218-1248 (row-wise), out [N,2d] */
};
/* gradient flavour for PNTF_FIELD_DIRECTION / pntf_plan / MODE_GRAD:
 * EXACT = true grad tau (S: autograd, NN.out_grad); COMPAT = what M:402-647 (out_backgrad) returns,
 * whose first encoder gate is evaluated at softplus(y) (M:435-438). */
enum { PNTF_GRAD_EXACT = 0, PNTF_GRAD_COMPAT = 1 };
/* workspace queries */
enum { PNTF_OP_FORWARD = 0, PNTF_OP_LOSS_FWD = 1, PNTF_OP_LOSS_BWD = 2, PNTF_OP_PLAN = 3, PNTF_OP_FIELD = 4 };

int pntf_version(void);
const char* pntf_last_error(void);

/* Offset / shape of tensor `index` (0..29, state_dict order: encoder.0.weight, encoder.0.bias, ...)
 * inside the flat parameter buffer.  Returns 0, or PNTF_ERR_ARG for a bad index. */
int pntf_param_offset(int index, int64_t* offset, int64_t* rows, int64_t* cols, int* used_by_forward);

/* Bytes of scratch the op needs for this shape (pass the same buffer as `workspace`).  For
 * PNTF_OP_LOSS_BWD `budget_bytes` (0 = default 8 GiB) caps the activation stash; larger batches are
 * processed in chunks inside one call. */
size_t pntf_workspace_bytes(int op, int E, int64_t N, int d, int prec, size_t budget_bytes);

/* NN.out / NN.out_grad / NN.out_laplace (S:212-256, 513-613, 676-820; M:215-259, 303-400, 710-848).
 * tau[E*N], dtau[E*N,2d] (mode>=GRAD), ltau[E*N,2d] (mode LAPLACE); unused outputs may be NULL. */
int pntf_forward(int mode, int grad_flavour, const float* params, const float* B, const float* coords,
                 int E, int64_t N, int d, float* tau, float* dtau, float* ltau, int prec,
                 void* workspace, size_t workspace_bytes, void* stream);

/* Model.TravelTimes / Speed / Speed2 / Gradient: out[E*N] (DIRECTION: out[E*N,2d]). */
int pntf_field(int kind, int grad_flavour, const float* params, const float* B, const float* coords,
               int E, int64_t N, int d, float gamma, float* out, int prec,
               void* workspace, size_t workspace_bytes, void* stream);

/* Model.Loss forward only (S:869-936, M:897-951).  speed[E*N,2] is used as alpha*speed+1-alpha
 * (the trainer's interpolation, S:1063 / M:1051; pass alpha = 1 for the raw Loss()).
 * diff[E*N] = per-pair residual; loss_out[2] = {loss = beta*loss_n, loss_n} (device). */
int pntf_loss_fwd(int variant, const float* params, const float* B, const float* coords, const float* speed,
                  int E, int64_t N, int d, float alpha, float beta, float gamma,
                  float* diff, float* loss_out, int prec,
                  void* workspace, size_t workspace_bytes, void* stream);

/* Model.Loss + loss.backward() (S:1065-1071, M:1070-1076): same outputs as pntf_loss_fwd, and
 * grad_params[PNTF_PARAM_COUNT] += d loss / d params (the caller zeroes it; encoder1.0.* is never
 * written -- the reference leaves its .grad None). */
int pntf_loss_bwd(int variant, const float* params, const float* B, const float* coords, const float* speed,
                  int E, int64_t N, int d, float alpha, float beta, float gamma,
                  float* diff, float* loss_out, float* grad_params, int prec,
                  void* workspace, size_t workspace_bytes, void* stream);

/* The same two calls with the batch gather of the training loop fused into the kernels' loads
 * (models/model_res_sigmoid_multi.py:1053-1066: `data[perm[lo:hi]]`, then points = batch[:, :2d],
 * speed = batch[:, 2d:]; single: FastTensorDataLoader, models/model_res_sigmoid.py:30-73).
 * dataset[dataset_rows, 2d+2] is the GPU-resident table in the reference's layout (sampled_points ++ speed,
 * models/data_multi.py:15-31; several environments are stacked row-wise); pair p of the [E, N] batch is row
 * index[p] of it.  Nothing is materialised: coordinates and speeds are read through the index.  Every index value
 * must lie in [0, dataset_rows): the kernels do not check it (the index lives on the device). */
int pntf_loss_fwd_gather(int variant, const float* params, const float* B, const float* dataset, int64_t dataset_rows,
                         const int32_t* index, int E, int64_t N, int d, float alpha, float beta, float gamma,
                         float* diff, float* loss_out, int prec,
                         void* workspace, size_t workspace_bytes, void* stream);
int pntf_loss_bwd_gather(int variant, const float* params, const float* B, const float* dataset, int64_t dataset_rows,
                         const int32_t* index, int E, int64_t N, int d, float alpha, float beta, float gamma,
                         float* diff, float* loss_out, float* grad_params, int prec,
                         void* workspace, size_t workspace_bytes, void* stream);

/* torch.optim.AdamW step on flat buffers (S:954-956,1074; M:960-962,1079): decoupled weight decay,
 * bias-corrected; elements whose mask byte is 0 are left untouched (parameters with grad None).
 * grad_scale multiplies the gradient first (1/world_size after an all-reduce(sum)).  Hyper-parameters
 * are doubles (Python floats): the derived constants 1-beta, lr/(1-beta1^t), ... are formed in double
 * exactly as torch does before they are rounded to fp32. */
int pntf_adamw(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, const uint8_t* mask,
               int64_t n, double lr, double beta1, double beta2, double eps, double weight_decay,
               int64_t step, double grad_scale, void* stream);

/* Batched bidirectional path integration (test/gib_plan.py:53-86, test/arm_plan.py:111-152):
 * per query, while |x1-x0| > thr: XP += eta*Gradient(XP), at most max_iter+1 steps.
 * waypoints[Q, max_iter+2, 2d]: row 0 = the query, row k = position after step k, rows after the
 * last step repeat the final position; nsteps[Q] = steps taken. */
int pntf_plan(const float* params, const float* B, const float* start_goal, int64_t Q, int d,
              float eta, float thr, int max_iter, int grad_flavour,
              float* waypoints, int32_t* nsteps, int prec,
              void* workspace, size_t workspace_bytes, void* stream);

/* Device-resident schedule (Model.train's epoch, M:1040-1097 / S:1050-1084, replayed as a CUDA graph).
 * The two calls of a training step with the per-epoch scalars read from DEVICE memory at run time, so that launches
 * captured once (cudaStreamBeginCapture ... EndCapture on `stream`) stay valid when the host changes the schedule
 * between replays:  sched[3] = {alpha, beta, gamma} (speed interpolation M:1051, loss weight M:1109, viscosity M:1014),
 * lr_step[2] = {learning rate M:1025, AdamW step count t as a float}.  Neither call allocates or synchronises. */
int pntf_loss_bwd_gather_sched(int variant, const float* params, const float* B, const float* dataset, int64_t dataset_rows,
                               const int32_t* index, int E, int64_t N, int d, const float* sched, float* diff,
                               float* loss_out, float* grad_params, int prec, void* workspace, size_t workspace_bytes,
                               void* stream);
int pntf_adamw_sched(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, const uint8_t* mask, int64_t n,
                     const float* lr_step, double beta1, double beta2, double eps, double weight_decay, double grad_scale,
                     void* stream);

/* Same integration; rows of waypoints after a query's last step (row nsteps[q]) are NOT written (the caller's buffer
 * keeps whatever it held).  pntf_plan fills them with the final position, which at the reference's cap of 500 is
 * 30x the bytes of the waypoints themselves; batched callers read waypoints[q, :nsteps[q]+1] only. */
int pntf_plan_compact(const float* params, const float* B, const float* start_goal, int64_t Q, int d,
                      float eta, float thr, int max_iter, int grad_flavour,
                      float* waypoints, int32_t* nsteps, int prec,
                      void* workspace, size_t workspace_bytes, void* stream);

/* Ground-truth speed sampling: unsigned distance from N query points [N,3] to a triangle soup
 * triangles[M,3,3] and, when speed != NULL, speed = clip(distance, offset, margin) / margin.
 * Replaces dataprocessing/speed_sampling_gpu.py:325-336 (point_obstacle_distance: sqrt of the squared distance of the
 * third-party bvh_distance_queries.BVH) and the clipping of :415-419 (point_rand_sample_bound_points).  Exact
 * point-triangle distances (closest point by Voronoi region), minimum over the mesh; results do not depend on the
 * order of triangles or points, spatially sorted inputs run faster.  Workspace: pntf_mesh_workspace_bytes(M). */
size_t pntf_mesh_workspace_bytes(int64_t M);
int pntf_mesh_distance(const float* triangles, int64_t M, const float* points, int64_t N, float* distance,
                       float offset, float margin, float* speed, void* workspace, size_t workspace_bytes,
                       void* stream);

/* Run-to-run reproducible gradients (tensor-core mode).  By default the weight-gradient and bias-gradient partial sums
 * of different CTAs / warps meet in fp32 atomics, whose order varies from run to run (last bits differ).  With on != 0
 * (or PNTF_DETERMINISTIC=1 in the environment) pntf_loss_bwd* accumulates them as 64-bit fixed point (2^-24) in its
 * workspace -- integer adds commute -- and converts once at the end: two calls on the same inputs return bit-identical
 * grad_params.  The reference's autograd on CUDA gives no such guarantee either; this is for debugging and for
 * replicas that must not drift.  Process-global. */
int pntf_set_deterministic(int on);

/* Measurement aid for bench.py: between begin and end every kernel (group) the library launches is
 * bracketed by CUDA events on its stream; end synchronises on them and returns, per kernel class
 * (0 pack, 1 forward, 2 reverse, 3 dW, 4 loss-finalize, 5 adamw, 6 plan, 7 tc-forward, 8 tc-reverse,
 * 9 tc-dW, 10 tc-backward = reverse sweep + weight gradients in one launch), the summed device time in ms and the
 * number of bracketed launches (groups).  Process-global, unsynchronised state: call from the one host thread that
 * drives the library, for measurement only. */
#define PNTF_KERNEL_CLASSES 11
int pntf_profile_begin(void);
int pntf_profile_end(double* ms_by_class, int64_t* count_by_class, int n_classes);

/* Diagnostic of the tcgen05 plumbing (operand layouts, descriptors, TMEM mapping); no reference
 * counterpart.  W, X, C: [128,128] fp32 device arrays; out[3,128,128] = {W X, C + W X, X X^T} computed
 * with fp16 operands and fp32 accumulation on the tensor cores. */
int pntf_tc_selftest(const float* W, const float* X, const float* C, float* out, void* stream);

/* Test aid for the tensor-core backward (no reference counterpart): statistics of the fp16 adjoint tiles the last
 * pntf_loss_bwd(prec = PNTF_PREC_TC) call left in `workspace` (same E, N, d; the call must have fitted one chunk).
 * Adjoints travel as scaled fp16; layer = -1: all adjoint tiles Ybar_0..12, 0..12: Ybar_layer, 100 + l: the forward
 * operand tile X_l (l = 0..13; a library built with -DPNTF_STASH_REGEN regenerates X_0 / X_6 instead of stashing them
 * and answers PNTF_ERR_ARG for those), 200: the stashed encoder outputs Y5.  out[4] (device, uint64) = {inf/NaN count, subnormal count, nonzero count, fp32 bit
 * pattern of the largest finite magnitude}. */
int pntf_debug_adjoint_stats(const void* workspace, size_t workspace_bytes, int E, int64_t N, int d, int layer,
                             uint64_t* out, void* stream);

/* Allocation-free path integration.  No counterpart in the reference: added for
 * callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).
 * Same computation and outputs as pntf_plan(prec = PNTF_PREC_TC) -- waypoints[Q, max_iter+2, 2d], nsteps[Q], both
 * caller-provided device buffers -- but the scratch (fp16 weight image, query counter, hand-out order) is a file-scope
 * __device__ array inside the library: 1 378 560 B (1.31 MiB) that every GPU which loads the library holds until the
 * process exits.  Limit: Q <= PNTF_PLAN_STATIC_MAX_QUERIES per call; a larger Q returns PNTF_ERR_WORKSPACE and launches
 * nothing.  The call does not allocate, free or synchronise (first call on a device: one cudaGetSymbolAddress).
 * All calls on a device share the one array: calls that overlap on the GPU (different streams without an ordering
 * between them) race on it -- keep them on one stream or order them with events.  Each call rewrites or clears
 * everything it reads, so nothing carries over from the previous call. */
#define PNTF_PLAN_STATIC_MAX_QUERIES 65536
int pntf_plan_static(const float* params, const float* B, const float* start_goal, int64_t Q, int d,
                     float eta, float thr, int max_iter, int grad_flavour,
                     float* waypoints, int32_t* nsteps, void* stream);

/* number of kernels the library has launched since load (for bench.py's gpu_launches) */
int64_t pntf_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* PNTF_B200_H_ */
