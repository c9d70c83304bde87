// tcgen05 forward kernel (tensor-core precision mode): a persistent, warp-specialised CTA carries
// tiles of P start/goal pairs through the whole network (Fourier features -> per-endpoint encoder ->
// symmetric combine -> generator -> output rules -> per-pair tail).  The 14 linear layers run as
// tcgen05.mma (fp16 operands, fp32 accumulators in TMEM); value / first-derivative / diagonal
// second-derivative columns are propagated in forward mode by the epilogue warps, which read the
// accumulators with tcgen05.ld, apply bias / residual skip / softplus rules and write the next
// layer's fp16 operand tile back to shared memory.  The residual skips stay in fp32: the block input
// is pre-loaded into the accumulator of the block's second layer with tcgen05.st.
//   warps 0..15 : epilogue; thread = TMEM lane (= output feature within a 128-feature half) x 32-row
//                 quarter (lane quarter = warp % 4, row quarter = warp / 4)
//   warp  16    : weight-stage producer (cp.async.bulk L2 -> shared memory ring)
//   warp  17    : MMA issuer
//   warp  18    : bulk stores of the training stash (TRAIN)
// Pipelining inside a tile: operand tiles and accumulators are handed over per 128-feature half
// (mbarriers xr[h] / ar[h]), so the issuer starts layer l+1 on input features 0..127 while the epilogue
// still produces features 128..255 of layer l, and the epilogue starts on output half 0 while half 1
// is still being multiplied.
// Operand tiles are double-buffered (X_l lives in buffer l & 1), so an epilogue never overwrites the tile
// the running product or stash store still reads.
// Training stash (TRAIN): the operand tile X_l of every layer (l = 0..13) is bulk-stored as is; the
// reverse kernel recovers the activation gates from X_{l+1} = act(Y_l) (s = 1 - exp(-10 a)), so the
// pre-activations are not stashed, except the encoder outputs Y_5 (input of the combine) and the fp32
// rows of the output layer.
// Reference: NN.out_laplace, models/model_res_sigmoid.py:676-820 (models/model_res_sigmoid_multi.py:710-848);
// NN.out_grad S:513-613; Model.Loss S:869-936 / M:897-951.
#pragma once
#include "math.cuh"
#include "pair_math.cuh"
#include "tc_layers.cuh"

namespace pntf {
namespace tc {

enum { TC_TAIL_NONE = 0, TC_TAIL_LOSS = 1, TC_TAIL_FIELD = 2, TC_TAIL_PLAN = 3 };

struct TcStash {
  uint8_t* X[kNumLayers];   // fp16 operand-tile images of the layer inputs, [tile][K_l * 256 B], l = 0..13
  uint8_t* Y5;              // fp16 tile images of the encoder outputs (pre-combine), [tile][128 * 256 B]
  uint8_t* G[kNumLayers];   // fp16 tile images of the pre-activation adjoints (reverse kernel), l = 0..12
  float* y13;               // [tile][128] fp32 pre-activation rows of the output layer
};

struct TcFwdArgs {
  const float* params;
  const uint8_t* wimg;      // forward stage images
  const float* B;           // [E][D][128]
  const float* coords;      // [n_total][2D]
  const float* speed;       // [n_total][2]
  const int32_t* index;     // optional batch gather (pair_math.cuh: src_row), else null
  int cstride, sstride;     // row strides of coords / speed in floats (2D / 2 when dense)
  int64_t n_total, n_per_env, pair_begin, pair_count;
  float* tau; float* dtau; float* ltau;
  float* field; int field_kind;
  float* diff; int single_variant;
  float alpha, gamma;
  const float* sched;       // optional device-resident {alpha, beta, gamma} (pair_math.cuh)
  int compat;
  TcStash st;
  long long* trace;         // debug (PNTF_TRACE=1): clock64 time stamps of CTA 0's third tile, else nullptr
  // TC_TAIL_PLAN (batched path integration, test/gib_plan.py:53-86): coords = start/goal queries
  float eta, thr;
  int max_iter;
  float* waypoints;         // [Q][max_iter+2][2D]
  int32_t* nsteps;          // [Q]
  int* tile_counter;        // zeroed by the host: path queries are handed out to free tile slots dynamically
  int fill_tail;            // 1: rows after a query's last step repeat the final position (pntf_plan); 0: left untouched
  const int32_t* order;     // hand-out order of the queries (longest start-goal distance first), or null: input order
  int coords_bulk;          // 1: dense, 16-byte aligned coordinates -> a tile's P x 2D floats arrive by one bulk copy (TMA)
};

// debug time stamps: slot layout [role][layer][event]
__device__ __forceinline__ void trace_evt(long long* trace, bool on, int role, int l, int evt) {
  if (trace != nullptr && on) trace[(role * 16 + l) * 4 + evt] = clock64();
}

template <class C>
struct TcTile {
  float x[C::P][2 * C::D];
  int env[C::P];
  int valid[C::P];
};

// softplus_10 with its derivative gates, fast intrinsics (TC mode is operand-rounded to fp16 anyway)
struct FastAct { float a, s, sp; };
__device__ __forceinline__ float ex2_approx(float x) {
  float r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ float lg2_approx(float x) {
  float r;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
// With z = 10 y:  e = exp(-|z|) = 2^(-10 log2(e) |y|),  softplus_10(y) = 0.1 (max(z, 0) + ln(1 + e)) = max(y, 0) + 0.1 ln(2) log2(1 + e):
// the scale factors are folded into the constants (three multiplications fewer per evaluation than the literal form; torch's
// threshold -- return y for z > 20 -- is what the formula gives in fp32 anyway).
__device__ __forceinline__ FastAct fast_act(float y) {
  FastAct r;
  const float e = ex2_approx(-14.4269504f * fabsf(y));
  const float inv = __fdividef(1.0f, 1.0f + e);
  r.s = (y >= 0.f) ? inv : e * inv;
  r.a = fmaf(lg2_approx(1.0f + e), 0.0693147181f, fmaxf(y, 0.f));
  r.sp = 10.0f * r.s * (1.0f - r.s);
  return r;
}
__device__ __forceinline__ float fast_sigmoid(float x) {
  const float e = __expf(-fabsf(x));
  const float inv = __fdividef(1.0f, 1.0f + e);
  return (x >= 0.f) ? inv : e * inv;
}

// hidden activation with the forward-mode rules on one unit (S:641-657): y[0] value, y[1..NDER] first derivatives
// (NDER = D for an encoder endpoint, 2D for a generator pair), then the second-order columns: one per coordinate
// (ORDER 2) or one sum per endpoint (kOrderLapSum: L <- s' sum_i dy_i^2 + s L).
template <int D, bool GEN, int ORDER>
__device__ __forceinline__ void act_unit(float* y, bool compat) {
  constexpr int NDER = ORDER >= 1 ? (GEN ? 2 * D : D) : 0;
  const FastAct a = fast_act(y[0]);
  y[0] = a.a;
  const float gate = compat ? fast_sigmoid(10.0f * a.a) : a.s;
  if constexpr (ORDER == kOrderLapSum) {
#pragma unroll
    for (int e = 0; e < (GEN ? 2 : 1); ++e) {
      float q = 0.f;
#pragma unroll
      for (int i = 0; i < D; ++i) q = fmaf(y[1 + e * D + i], y[1 + e * D + i], q);
      y[1 + NDER + e] = fmaf(q * kLapScale, a.sp, y[1 + NDER + e] * a.s);
    }
  }
#pragma unroll
  for (int i = 0; i < NDER; ++i) {
    const float dy = y[1 + i];
    if constexpr (ORDER == 2) y[1 + NDER + i] = fmaf(dy * dy * kLapScale, a.sp, y[1 + NDER + i] * a.s);
    if constexpr (ORDER >= 1) y[1 + i] = dy * gate;
  }
}

// Position of an epilogue thread: TMEM lane (feature within a half), 32-row quarter, and the hoisted
// parts of the operand-tile addressing for its rows.
struct EpiPos {
  int lane;         // 0..127
  int rq;           // row quarter 0..3 (rows 32*rq .. 32*rq+31)
  int nb, cb;       // 64-row block and first 8-row chunk inside it
  uint32_t tmem;    // TMEM address of (lane quarter, column 0)
};
__device__ __forceinline__ EpiPos epi_pos(uint32_t tmem_base) {
  EpiPos p;
  const int warp = threadIdx.x >> 5;
  p.lane = (warp & 3) * 32 + (threadIdx.x & 31);
  p.rq = warp >> 2;
  p.nb = p.rq >> 1;
  p.cb = (p.rq & 1) * 4;
  p.tmem = tmem_base + (uint32_t((warp & 3) * 32) << 16);
  return p;
}

// store the thread's 32 rows of feature f as fp16 (4 chunks)
__device__ __forceinline__ void st_rows32(uint32_t act_addr, const EpiPos& p, int f, const float* v) {
  const uint32_t line = act_line_addr(act_addr, f, p.nb);
#pragma unroll
  for (int j = 0; j < 4; ++j) sts_chunk(line, f & 7, p.cb + j, v + 8 * j);
}
// load the thread's 32 rows of feature f from a tile image in shared memory
__device__ __forceinline__ void ld_rows32(uint32_t tile_addr, const EpiPos& p, int f, float* v) {
  const uint32_t line = act_line_addr(tile_addr, f, p.nb);
#pragma unroll
  for (int j = 0; j < 4; ++j) lds_chunk(line, f & 7, p.cb + j, v + 8 * j);
}
// Epilogue of one 128-feature half of a hidden layer for the thread's 32 rows.
//   ENC: a pair's CGP columns are two endpoint units of CEP columns (D derivative columns each),
//   otherwise one unit of CGP columns (2D derivative columns).
//   PRELOAD: also store the fp32 result to accumulator pre_col (residual skip of the next block).
template <class C, bool ENC, bool PRELOAD>
__device__ __forceinline__ void epi_hidden_half(const EpiPos& p, int acc_col, int pre_col, int f, float b, uint32_t act_addr,
                                                bool compat) {
  constexpr int ST = C::STRIDE, EO = C::EOFF;
  float v[32];
  tmem_ld<32>(p.tmem + acc_col + 32 * p.rq, v);
#pragma unroll
  for (int u = 0; u < C::UPT; ++u) {
    float* y = v + u * ST;
    if constexpr (ENC) {
      y[0] += b;
      y[EO] += b;
      act_unit<C::D, false, C::ORDER>(y, compat);
      act_unit<C::D, false, C::ORDER>(y + EO, compat);
    } else {
      y[0] += b;
      act_unit<C::D, true, C::ORDER>(y, false);
    }
  }
  st_rows32(act_addr, p, f, v);
  if constexpr (PRELOAD) {
    tmem_st<32>(p.tmem + pre_col + 32 * p.rq, v);
    tmem_st_wait();
  }
}

// symmetric combine of one pair (S:724-787): u, w = the two endpoint encodings of a feature (CEP columns
// each); writes the smooth-max (MAXSIDE) or smooth-min output columns (CGP) to o.
template <class C, bool MAXSIDE>
__device__ __forceinline__ void combine_unit(const float* u, const float* w, float* o) {
  constexpr int D = C::D;
#pragma unroll
  for (int c = 0; c < C::STRIDE; ++c) o[c] = 0.f;
  const float dlt = u[0] - w[0];
  const float e = __expf(-10.0f * fabsf(dlt));
  const float tail = 0.1f * __logf(1.0f + e);
  o[0] = MAXSIDE ? fmaxf(u[0], w[0]) + tail : fminf(u[0], w[0]) - tail;
  if constexpr (C::ORDER >= 1) {
    const float inv = __fdividef(1.0f, 1.0f + e);
    const float s0 = (dlt >= 0.f) ? inv : e * inv;
    const float s1 = 1.0f - s0;
    const float q = (MAXSIDE ? 10.0f * s0 * s1 : -10.0f * s0 * s1) * kLapScale;   // second-order columns are scaled (tc_layers.cuh)
    const float ga = MAXSIDE ? s0 : s1, gb = MAXSIDE ? s1 : s0;   // gates of the start / goal side
    float qa = 0.f, qb = 0.f;
#pragma unroll
    for (int i = 0; i < D; ++i) {
      const float da = u[1 + i], db = w[1 + i];
      o[1 + i] = da * ga;
      o[1 + D + i] = db * gb;
      if constexpr (C::ORDER == 2) {
        o[1 + 2 * D + i] = fmaf(q * da, da, u[1 + D + i] * ga);
        o[1 + 3 * D + i] = fmaf(q * db, db, w[1 + D + i] * gb);
      }
      qa = fmaf(da, da, qa);
      qb = fmaf(db, db, qb);
    }
    if constexpr (C::LS) {   // summed second-order columns: start side, goal side
      o[1 + 2 * D] = fmaf(q, qa, u[1 + D] * ga);
      o[1 + 2 * D + 1] = fmaf(q, qb, w[1 + D] * gb);
    }
  }
}

// Epilogue of encoder layer 5 (no activation) fused with the symmetric combine: reads the two endpoint
// encodings of feature `lane`, writes smooth-max (feature lane) and smooth-min (feature 128+lane) with
// their derivative columns, and pre-loads both into the skip accumulator.  y5_addr != 0: also writes the
// encoder outputs (bias included) to a 128-feature staging image in shared memory (training stash).
template <class C>
__device__ __forceinline__ void epi_combine(const EpiPos& p, int acc_col, int pre_col, float b, uint32_t act_addr, uint32_t y5_addr) {
  constexpr int ST = C::STRIDE, EO = C::EOFF, UPT = C::UPT;
  float v[32], o[32];
  tmem_ld<32>(p.tmem + acc_col + 32 * p.rq, v);
#pragma unroll
  for (int u = 0; u < UPT; ++u) {
    v[u * ST] += b;
    v[u * ST + EO] += b;
  }
  if constexpr (UPT * ST < 32) {
#pragma unroll
    for (int c = UPT * ST; c < 32; ++c) o[c] = 0.f;   // unused rows of the quarter stay zero
  }
  if (y5_addr != 0) st_rows32(y5_addr, p, p.lane, v);
#pragma unroll
  for (int u = 0; u < UPT; ++u) combine_unit<C, true>(v + u * ST, v + u * ST + EO, o + u * ST);
  st_rows32(act_addr, p, p.lane, o);
  tmem_st<32>(p.tmem + pre_col + 32 * p.rq, o);
#pragma unroll
  for (int u = 0; u < UPT; ++u) combine_unit<C, false>(v + u * ST, v + u * ST + EO, o + u * ST);
  st_rows32(act_addr, p, 128 + p.lane, o);
  tmem_st<32>(p.tmem + pre_col + 128 + 32 * p.rq, o);
  tmem_st_wait();
}

// Fourier features of both endpoints of the thread's pairs (S:204-210, M:199-213) into the layer-0 operand tile.
// One call = (frequency k, row quarter rq): features k (sin) and 128 + k (cos), 32 rows.  Also used by the weight-gradient
// kernel, which regenerates this tile instead of reading it from the stash (tc_dw.cuh).
template <class C>
__device__ __forceinline__ void fourier_rows(const float* __restrict__ Bmat, const TcTile<C>& ts, int kfreq, int rq, uint32_t act_addr) {
  constexpr int ST = C::STRIDE, EO = C::EOFF, D = C::D;
  EpiPos p;
  p.lane = kfreq;
  p.rq = rq;
  p.nb = p.rq >> 1;
  p.cb = (p.rq & 1) * 4;
  p.tmem = 0;
  const int k = p.lane;
  float vs[32], vc[32];
#pragma unroll
  for (int c = 0; c < 32; ++c) { vs[c] = 0.f; vc[c] = 0.f; }
#pragma unroll
  for (int u = 0; u < C::UPT; ++u) {
    const int pr = p.rq * C::UPT + u;
    const float* Bp = Bmat + int64_t(ts.env[pr]) * D * kNumFreq + k;
    float w[D];
#pragma unroll
    for (int i = 0; i < D; ++i) w[i] = 6.2831855f * __ldg(Bp + i * kNumFreq);
#pragma unroll
    for (int side = 0; side < 2; ++side) {
      float proj = 0.f;
#pragma unroll
      for (int i = 0; i < D; ++i) proj = fmaf(ts.x[pr][side * D + i], w[i], proj);
      float sn, cs;
      // one-step range reduction to [-pi, pi], then the fast intrinsics (abs. error ~1e-6; the tile is fp16)
      const float red = fmaf(-rintf(proj * 0.15915494f), 6.2831855f, proj);
      __sincosf(red, &sn, &cs);
      float* us = vs + u * ST + side * EO;
      float* uc = vc + u * ST + side * EO;
      us[0] = sn;
      uc[0] = cs;
      if constexpr (C::ORDER >= 1) {
        float w2 = 0.f;
#pragma unroll
        for (int i = 0; i < D; ++i) {
          us[1 + i] = w[i] * cs;
          uc[1 + i] = -w[i] * sn;
          if constexpr (C::ORDER == 2) {
            us[1 + D + i] = -w[i] * w[i] * kLapScale * sn;
            uc[1 + D + i] = -w[i] * w[i] * kLapScale * cs;
          }
          w2 = fmaf(w[i], w[i], w2);
        }
        if constexpr (C::LS) {
          us[1 + D] = -w2 * kLapScale * sn;
          uc[1 + D] = -w2 * kLapScale * cs;
        }
      }
    }
  }
  st_rows32(act_addr, p, k, vs);
  st_rows32(act_addr, p, 128 + k, vc);
}
// forward kernels: thread = (frequency, row quarter) over the 512 epilogue threads
template <class C>
__device__ __forceinline__ void fourier_tile(const TcFwdArgs& a, const TcTile<C>& ts, uint32_t act_addr) {
  fourier_rows<C>(a.B, ts, int(threadIdx.x & 127), int(threadIdx.x >> 7), act_addr);
}

// shared-memory control block
template <int S>
struct TcCtl {
  uint64_t full[S], empty[S];   // weight ring
  uint64_t xr[2];               // operand-tile half h complete (all epilogue threads arrive)
  uint64_t ar[2];               // accumulator half h complete (tcgen05.commit)
  uint64_t xio[2][2];           // TRAIN: operand-tile half h complete, as forwarded by the MMA issuer to the stash-store
                                // warp; successive events alternate between two barriers (see the issuer loop)
  uint64_t sd[2][2];            // stash store of a buffer half h has finished reading shared memory; successive
                                // events alternate between two barriers (a waiter may lag two events behind)
  uint32_t tmem_base;
  // TC_TAIL_PLAN: the number of passes is data dependent; an epilogue thread announces every pass
  uint64_t go;                  // completes once per announcement
  volatile int more;            // 1 = another pass follows, 0 = the CTA is done
  uint32_t ready;               // PNTF_HELPER_POLL: weight stages landed so far (written by the helper thread)
  uint64_t cbar[2];             // coordinates of the tile in ts[i] have landed (one completion per load_tile call)
};

// all epilogue threads (converged warps): half h of the next operand tile (and its TMEM pre-loads) is complete.
// Every thread fences its own writes; one lane per warp arrives (the barrier counts warps), which keeps the
// mbarrier unit free for the MMA issuer's polls.
// Publishing an operand-tile half.  Forward kernel: every thread fences its own writes, __syncwarp(), one lane per
// warp arrives (the barrier counts warps): 3.5 ms per 1 Mi pairs faster than 512 per-thread arrivals, which compete
// with the MMA issuer's polls on the mbarrier unit.  (With the former stash-store protocol this faster publish exposed
// a dead-lock on cold first launches; see the issuer loop.  scripts/stress.sh: 0 failures of 60 since.)
// Reverse kernel: every thread arrives by itself (measured faster there: longer, more skewed epilogues).
constexpr int kPubCount = kEpiWarps;
__device__ __forceinline__ void epi_publish(uint64_t* bar) {
  fence_proxy_async();
  tc_fence_before();
  __syncwarp();
  if ((threadIdx.x & 31) == 0) mbar_arrive(bar);
}
__device__ __forceinline__ void epi_publish_each(uint64_t* bar) {
  fence_proxy_async();
  tc_fence_before();
  mbar_arrive(bar);
}

// MMA issue of one (contraction half kh, output half h) block: 2 stages x 4 UMMAs.  Called by the whole
// (converged) MMA warp; one elected lane issues.
// Measured (scripts/umma_bench.cu, profiles/r01c_umma_microbench.md): a poll of a ready mbarrier costs the issuing
// warp 100-250 clk and the UMMA queue is shallow, so the per-stage poll + commit keeps the pipe at ~110-130 clk per
// UMMA (floor 64).  A ring of three 32 KB blocks (one poll per 8 UMMAs, same smem look-ahead) was tried and was
// 4 ms per 1 Mi pairs SLOWER in the training forward (coarser slot release), so the 16 KB stages stay.
// Weight-ring slot release.  Default: one tcgen05.commit per 16 KB stage.  -DPNTF_PAIR_COMMIT (experiment): mma_block
// consumes stages in pairs (2m, 2m+1) and commits once, after the second, on that stage's barrier; the producer, before
// refilling the slot of stage n (last used by stage c = n - S), waits for the commit of the pair c belongs to.
template <int S, class CTL>
__device__ __forceinline__ void ring_wait_free(CTL* ctl, uint32_t n) {
#ifdef PNTF_PAIR_COMMIT
  if (n < uint32_t(S)) return;
  const uint32_t c = (n - S) | 1u;   // the odd stage of the pair
  mbar_wait(&ctl->empty[c % S], (c / (2 * S)) & 1);
#else
  mbar_wait(&ctl->empty[n % S], ((n / S) & 1) ^ 1);
#endif
}
#ifndef PNTF_NO_LOOKAHEAD
#define PNTF_LOOKAHEAD 1   // +4-5 % on the no-stash forward (profiles/r01m_*); -DPNTF_NO_LOOKAHEAD restores the plain poll
#endif
template <int S, bool HELPER = false, class CTL>
__device__ __forceinline__ void mma_block(CTL* ctl, uint32_t ring_addr, uint32_t opnd_addr, uint32_t d_tmem, int kh,
                                          bool accumulate_first, uint32_t& it) {
  constexpr uint32_t idesc = make_idesc(128, kRows, false, true);
  constexpr uint32_t a_hi = desc_hi(1024), b_hi = desc_hi(kActSBO);
#pragma unroll
  for (int s2 = 0; s2 < 2; ++s2) {
#if defined(PNTF_HELPER_POLL)
    // experiment: a helper thread observes the `full` barriers and publishes the number of landed stages in shared
    // memory; the issuer reads that word (a shared-memory load instead of a barrier poll)
    uint32_t slot;
    if constexpr (HELPER) {
      slot = it % S;
      uint32_t spins = 0;
      while (ld_acquire_u32(&ctl->ready) <= it) { if (++spins > (1u << 30)) __trap(); }
    } else {
      slot = it % S;
      mbar_spin(&ctl->full[slot], (it / S) & 1);
    }
#elif defined(PNTF_LOOKAHEAD)
    // Bit 31 of `it`: this stage was already seen full by the look-ahead poll of the previous stage.  The poll of the
    // NEXT stage's barrier is issued before this stage's UMMAs and its result is read after them, so its latency
    // (100-250 clk for a ready barrier) is hidden behind the issue instead of preceding it.
    const uint32_t n = it & 0x7fffffffu;
    const uint32_t slot = n % S, ph = (n / S) & 1;
    if (!(it >> 31)) mbar_spin(&ctl->full[slot], ph);
    __syncwarp();
    uint32_t ok_next;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok_next)
        : "r"(smem_u32(&ctl->full[(n + 1) % S])), "r"(((n + 1) / S) & 1)
        : "memory");
#else
    const uint32_t slot = it % S, ph = (it / S) & 1;
    mbar_spin(&ctl->full[slot], ph);
#endif
    tc_fence_after();
    if (elect_one()) {
      const uint32_t a_lo = desc_lo(ring_addr + slot * kStageBytes, 16);
      const uint32_t b_lo = desc_lo(opnd_addr + (kh * 2 + s2) * 8 * kActSBO, kActLBO);
#pragma unroll
      for (int j = 0; j < 4; ++j)
        umma_f16_split(d_tmem, a_lo + 2 * j, a_hi, b_lo + j * ((2 * kActSBO) >> 4), b_hi, idesc,
                       (accumulate_first || s2 > 0 || j > 0) ? 1u : 0u);
#ifdef PNTF_PAIR_COMMIT
      if (s2 == 1) umma_commit(&ctl->empty[slot]);
#else
      umma_commit(&ctl->empty[slot]);
#endif
    }
    __syncwarp();
#if defined(PNTF_HELPER_POLL)
    ++it;
#elif defined(PNTF_LOOKAHEAD)
    it = (n + 1) | (ok_next << 31);
#else
    ++it;
#endif
  }
}

// One thread issues a 16 KB bulk copy every ~230-260 clk (64-73 B/clk, scripts/l2_stream.cu, scripts/umma_smem_bw.cu):
// exactly what the tensor pipe consumes at its peak, with no slack.  Two lanes of the producer warp, taking the
// stages in turn, deliver ~100 B/clk without another warp (more warps made ptxas spill).
#ifndef PNTF_PRODUCER_LANES
#define PNTF_PRODUCER_LANES 1
#endif
constexpr int kProducerLanes = PNTF_PRODUCER_LANES;
constexpr int kFwdRing = 5;                 // weight ring depth of the forward kernel

// Second MMA issuer (kernels without a stash, whose I/O warp is otherwise idle): in the 256-feature layers the MMA warp
// issues output half 0 and the I/O warp output half 1, each on its own accumulator columns and with its own commits; both
// take their weight stages from the one ring in image order (the MMA warp skips the stages of half 1 and vice versa).
// In isolation two issuers reach 79 clk per UMMA against 105-113 for one (profiles/r01c_umma_microbench.md, section 6).
#ifndef PNTF_DUAL_ISSUER
#define PNTF_DUAL_ISSUER 0
#endif

template <int D, int ORDER, bool TRAIN, int TAIL>
__global__ void __launch_bounds__(kTcThreads, 1) tc_forward_kernel(const TcFwdArgs a) {
  using C = TcCfg<D, ORDER>;
  constexpr int S = kFwdRing;
  constexpr bool DUAL = (PNTF_DUAL_ISSUER != 0) && !TRAIN;
  constexpr int CGP = C::CGP, P = C::P;
  constexpr bool PLAN = (TAIL == TC_TAIL_PLAN);
  // who fetches a tile's coordinates: the MMA warp with one bulk copy per tile (COORD_TMA), or -- value-only rows, 64 pairs
  // per tile, where the extra work in the MMA warp measured 12 % slower -- the tail threads with per-thread loads
  constexpr bool COORD_TMA = !PLAN && ORDER != 0;
  static_assert(!PLAN || (ORDER == 1 && !TRAIN), "path integration runs the first-order forward");
  constexpr int ACC_A = 0, ACC_B = 256;       // TMEM accumulator regions (columns)
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* act = smem;                        // two operand-tile buffers: X_l in buffer l & 1
  uint8_t* ring = smem + 2 * kTileBytes;
  __shared__ TcCtl<S> ctl;
  __shared__ __align__(16) TcTile<C> ts[2];
  __shared__ float y13s[2][kRows];

  const int tid = threadIdx.x, warp = tid >> 5, lane_id = tid & 31;
  if (tid == 0) {
    for (int i = 0; i < S; ++i) { mbar_init(&ctl.full[i], 1); mbar_init(&ctl.empty[i], 1); }
    mbar_init(&ctl.xr[0], kPubCount);
    mbar_init(&ctl.xr[1], kPubCount);
    mbar_init(&ctl.ar[0], 1);
    mbar_init(&ctl.ar[1], 1);
    for (int i = 0; i < 4; ++i) mbar_init(&ctl.sd[i >> 1][i & 1], 1);
    for (int i = 0; i < 4; ++i) mbar_init(&ctl.xio[i >> 1][i & 1], 1);
    mbar_init(&ctl.go, 1);
    mbar_init(&ctl.cbar[0], 1);
    mbar_init(&ctl.cbar[1], 1);
    ctl.ready = 0;
    fence_barrier_init();
  }
  if (warp == kMmaWarp) tmem_alloc(&ctl.tmem_base, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = ctl.tmem_base;
  const int64_t ntiles = (a.pair_count + P - 1) / P;

  if (warp == kProducerWarp) {
    // ------------------------------------------------------------- weight-stage producer
    if (lane_id < kProducerLanes) {
      uint32_t it = 0, npass = 0;
      for (int64_t tile = blockIdx.x; PLAN || tile < ntiles; tile += gridDim.x) {
        if constexpr (PLAN) {
          mbar_wait(&ctl.go, npass & 1);
          ++npass;
          if (!ctl.more) break;
        }
        for (int g = 0; g < kFwdStages; ++g, ++it) {
          if (int(it % kProducerLanes) != lane_id) continue;   // the lanes take the stages in turn
          const uint32_t slot = it % S;
          ring_wait_free<S>(&ctl, it);
          mbar_arrive_expect_tx(&ctl.full[slot], kStageBytes);
          bulk_g2s(ring + slot * kStageBytes, a.wimg + size_t(g) * kStageBytes, kStageBytes, &ctl.full[slot]);
        }
      }
    }
  } else if (warp == kMmaWarp) {
    // ------------------------------------------------------------- MMA issuer (converged warp, elected lane issues)
    uint32_t it = 0, nx0 = 0, nx1 = 0, npass = 0;
    const uint32_t ring_addr = smem_u32(ring), act_addr = smem_u32(act);
    // Coordinates of tile `tile` into ts[i] (north_star: "streams coordinate pairs into shared memory with TMA").  Dense
    // input (coords_bulk) and a full tile: the tile's P x 2D floats are contiguous in global memory and in ts[i].x, ONE
    // bulk copy (cp.async.bulk) fetches them; ragged last tile / gathered batches (pntf_loss_*_gather): per-lane loads.
    // cbar[i] completes once per call; the epilogue threads wait on it before the tile's Fourier features.  This warp
    // does it because its register budget is free (the same fetch in the epilogue warps made ptxas spill there) and
    // because it knows when ts[i] is free: once it has seen the first operand half of tile k + 1, every epilogue warp
    // has left tile k's tail, the last reader of ts[k & 1].
    static_assert(sizeof(TcTile<C>) % 16 == 0 && (P * 2 * D * sizeof(float)) % 16 == 0, "bulk copies move 16-byte units");
    auto load_coords = [&](int64_t tile, int i) {
      TcTile<C>& t = ts[i];
      const int64_t first = a.pair_begin + tile * P;
      const int64_t end = (a.pair_begin + a.pair_count < a.n_total) ? a.pair_begin + a.pair_count : a.n_total;
      const bool bulk = a.coords_bulk != 0 && tile < ntiles && first + P <= end;
#pragma unroll
      for (int base = 0; base < P; base += 32) {
        const int k = base + lane_id;
        if (k < P) {
          const int64_t pair = first + k;
          const bool valid = tile < ntiles && pair < end;
          t.valid[k] = valid;
          t.env[k] = valid ? int(pair / a.n_per_env) : 0;
          if (!bulk) {
#pragma unroll
            for (int c = 0; c < 2 * D; ++c) t.x[k][c] = valid ? __ldg(a.coords + src_row(a, pair) * a.cstride + c) : 0.f;
          }
        }
      }
      __syncwarp();
      if (lane_id == 0) {
        if (bulk) {
          fence_proxy_async();   // earlier generic-proxy reads of ts[i].x are ordered before the async-proxy write
          mbar_arrive_expect_tx(&ctl.cbar[i], uint32_t(P * 2 * D * sizeof(float)));
          bulk_g2s(&t.x[0][0], a.coords + first * 2 * D, uint32_t(P * 2 * D * sizeof(float)), &ctl.cbar[i]);
        } else {
          mbar_arrive(&ctl.cbar[i]);
        }
      }
      __syncwarp();
    };
    if constexpr (COORD_TMA) {
      load_coords(blockIdx.x, 0);
      load_coords(int64_t(blockIdx.x) + gridDim.x, 1);
    }
    uint32_t ktile = 0;
    for (int64_t tile = blockIdx.x; PLAN || tile < ntiles; tile += gridDim.x, ++ktile) {
      if constexpr (PLAN) {
        mbar_wait(&ctl.go, npass & 1);
        ++npass;
        if (!ctl.more) break;
      }
#pragma unroll 1
      for (int l = 0; l < kNumLayers; ++l) {
        const int K = c_K[l], KH = K >> 7, MH = halves(c_M[l]);
        const bool skip = c_fwd_skip[l] != 0;   // accumulate onto the pre-loaded residual skip
        const uint32_t acc = tmem_base + (skip ? ACC_B : ACC_A);
#pragma unroll 1
        for (int kh = 0; kh < KH; ++kh) {
          // The operand tiles are double-buffered, so the epilogue may publish X_{l+1} (other buffer) before the
          // stash-store warp has looked at X_l: that warp can fall two events behind xr[kh], which a parity wait
          // cannot tell from zero events behind (observed as a dead-lock on cold first launches, when the first
          // stores into a fresh stash are slow).  The issuer sees every event in order (it is on the critical
          // path), so it forwards them to the store warp on two alternating barriers, each of which is then at
          // most one phase ahead of its waiter.
          if (kh == 0) {
            mbar_spin(&ctl.xr[0], nx0 & 1);
            if constexpr (TRAIN) { if (lane_id == 0) mbar_arrive(&ctl.xio[0][nx0 & 1]); __syncwarp(); }
            ++nx0;
          } else {
            mbar_spin(&ctl.xr[1], nx1 & 1);
            if constexpr (TRAIN) { if (lane_id == 0) mbar_arrive(&ctl.xio[1][nx1 & 1]); __syncwarp(); }
            ++nx1;
          }
          tc_fence_after();
          trace_evt(a.trace, lane_id == 0 && blockIdx.x == 0 && tile == int64_t(blockIdx.x) + 2 * gridDim.x, 0, l, kh);   // operand half kh seen
#pragma unroll 1
          for (int h = 0; h < MH; ++h) {
            if (DUAL && h == 1) {       // the second issuer's block: skip its two stages (and drop the look-ahead flag, which
              it = (it & 0x7fffffffu) + 2;   // would refer to one of them)
              continue;
            }
            mma_block<S, !TRAIN>(&ctl, ring_addr, act_addr + (l & 1) * kTileBytes, acc + h * 128, kh, skip || kh > 0, it);
            if (kh == KH - 1) {
              if (elect_one()) umma_commit(&ctl.ar[h]);
              __syncwarp();
              trace_evt(a.trace, lane_id == 0 && blockIdx.x == 0 && tile == int64_t(blockIdx.x) + 2 * gridDim.x, 0, l, 2 + h);   // commit of half h issued
            }
          }
        }
        if constexpr (COORD_TMA) {
          // layer 0 of tile k is issued: tile k - 1's tail is over, its coordinate buffer takes tile k + 1's pairs
          if (l == 0 && ktile > 0) load_coords(tile + int64_t(gridDim.x), int((ktile + 1) & 1));
        }
      }
    }
  } else if (warp == kIoWarp) {
    if constexpr (DUAL) {
      // ----------------------------------------------------------- second MMA issuer: output half 1 of the 256-feature layers
      // It follows every operand event (also those of the 128-feature layers it has nothing to issue for) so that its
      // parity counters stay in step; in the layers it issues for, the epilogue cannot publish the next operand tile
      // before this warp's commit on ar[1], so it is never more than one phase behind xr[].
      uint32_t it = 0, nx0 = 0, nx1 = 0, npass = 0;
      const uint32_t ring_addr = smem_u32(ring), act_addr = smem_u32(act);
      for (int64_t tile = blockIdx.x; PLAN || tile < ntiles; tile += gridDim.x) {
        if constexpr (PLAN) {
          mbar_wait(&ctl.go, npass & 1);
          ++npass;
          if (!ctl.more) break;
        }
#pragma unroll 1
        for (int l = 0; l < kNumLayers; ++l) {
          const int KH = c_K[l] >> 7, MH = halves(c_M[l]);
          const bool skip = c_fwd_skip[l] != 0;
          const uint32_t acc = tmem_base + (skip ? ACC_B : ACC_A) + 128;
#pragma unroll 1
          for (int kh = 0; kh < KH; ++kh) {
            if (kh == 0) { mbar_spin(&ctl.xr[0], nx0 & 1); ++nx0; }
            else { mbar_spin(&ctl.xr[1], nx1 & 1); ++nx1; }
            if (MH == 1) { it += 2; continue; }
            tc_fence_after();
            it += 2;                    // the first issuer's block (half 0)
            mma_block<S, false>(&ctl, ring_addr, act_addr + (l & 1) * kTileBytes, acc, kh, skip || kh > 0, it);
            it &= 0x7fffffffu;          // the look-ahead flag refers to a stage of the first issuer
            if (kh == KH - 1) {
              if (elect_one()) umma_commit(&ctl.ar[1]);
              __syncwarp();
            }
          }
        }
      }
    }
#ifdef PNTF_HELPER_POLL
    if constexpr (!TRAIN) {
      // helper: follows the producer's stage sequence, observes each `full` barrier and publishes the stage count
      if (lane_id == 0) {
        uint32_t it = 0, npass = 0;
        for (int64_t tile = blockIdx.x; PLAN || tile < ntiles; tile += gridDim.x) {
          if constexpr (PLAN) {
            mbar_wait(&ctl.go, npass & 1);
            ++npass;
            if (!ctl.more) break;
          }
          for (int g = 0; g < kFwdStages; ++g) {
            mbar_spin(&ctl.full[it % S], (it / S) & 1);
            ++it;
            st_release_u32(&ctl.ready, it);
          }
        }
      }
    }
#endif
    // ------------------------------------------------------------- training-stash stores
    if constexpr (TRAIN) {
      if (lane_id == 0) {
        uint32_t nx0 = 0, nx1 = 0, ne[2] = {0, 0};
        for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
#pragma unroll 1
          for (int l = 0; l < kNumLayers; ++l) {
            const int K = c_K[l], KH = K >> 7;
#pragma unroll 1
            for (int kh = 0; kh < KH; ++kh) {
#ifdef PNTF_EXP_IODELAY     // test hook (scripts/build_variants.py x:-DPNTF_EXP_IODELAY=6000): a slow store warp, as on a cold first launch
              __nanosleep(PNTF_EXP_IODELAY);
#endif
              if (kh == 0) { mbar_wait(&ctl.xio[0][nx0 & 1], (nx0 >> 1) & 1); ++nx0; }
              else { mbar_wait(&ctl.xio[1][nx1 & 1], (nx1 >> 1) & 1); ++nx1; }
              if (l == 6 && kh == 0) {
                // encoder outputs Y5, staged by the combine epilogue in the upper half of buffer 1
                stash_s2g(a.st.Y5 + size_t(tile) * kHalfTileBytes, act + kTileBytes + kHalfTileBytes, kHalfTileBytes);
                bulk_commit();
                bulk_wait_read<0>();
                mbar_arrive(&ctl.sd[1][ne[1] & 1]);
                ++ne[1];
              }
              if (!stash_regen_layer(l)) {   // X_0 and X_6 are regenerated by the weight-gradient kernel (tc_layers.cuh)
                stash_s2g(a.st.X[l] + size_t(tile) * K * 256 + size_t(kh) * kHalfTileBytes,
                          act + (l & 1) * kTileBytes + kh * kHalfTileBytes, kHalfTileBytes);
                bulk_commit();
                bulk_wait_read<0>();
              }
              mbar_arrive(&ctl.sd[kh][ne[kh] & 1]);    // this half of the buffer may be overwritten (by the epilogue of layer l+1)
              ++ne[kh];
            }
          }
        }
        bulk_wait0();
      }
    }
  } else {
    // ------------------------------------------------------------- epilogue warps
    const EpiPos p = epi_pos(tmem_base);
    const uint32_t act_addr = smem_u32(act);
    const float* prm = a.params;
    uint32_t nar0 = 0, nar1 = 0;
    const bool compat = a.compat != 0;
    // coordinates of the CTA's first two tiles (later tiles are fetched two tiles ahead by the tail threads)
    // COORD_TMA: the coordinates of a tile (ts[k & 1]) are fetched by the MMA warp (load_coords above); cbar[k & 1]
    // completes when they have landed.  Otherwise the tail threads fetch them two tiles ahead with per-thread loads.
    uint32_t ntile = 0;      // tiles this CTA has started: tile k uses ts[k & 1], whose barrier is then in phase (k >> 1) & 1
    auto load_tile = [&](int64_t tile, TcTile<C>& t) {
      const int64_t pair = a.pair_begin + tile * P + tid;
      const bool valid = tile < ntiles && pair < a.pair_begin + a.pair_count && pair < a.n_total;
      t.valid[tid] = valid;
      t.env[tid] = valid ? int(pair / a.n_per_env) : 0;
#pragma unroll
      for (int i = 0; i < 2 * D; ++i) t.x[tid][i] = valid ? __ldg(a.coords + src_row(a, pair) * a.cstride + i) : 0.f;
    };
    if constexpr (!PLAN && !COORD_TMA) {
      if (tid < P) {
        load_tile(blockIdx.x, ts[0]);
        load_tile(int64_t(blockIdx.x) + gridDim.x, ts[1]);
      }
      named_bar_sync(1, kEpiThreads);
    }
    int par = 0;
    uint32_t nsd0 = 0, nsd1 = 0;
    const uint32_t buf0 = act_addr, buf1 = act_addr + kTileBytes;
    // TRAIN: before a buffer half is overwritten the stash store of its previous occupant (the operand tile of two
    // layers earlier) must have finished reading it; sd[h] completions are consumed strictly in order.
#define PNTF_WAIT_SD0() { if constexpr (TRAIN) { mbar_wait(&ctl.sd[0][nsd0 & 1], (nsd0 >> 1) & 1); ++nsd0; } }
#define PNTF_WAIT_SD1() { if constexpr (TRAIN) { mbar_wait(&ctl.sd[1][nsd1 & 1], (nsd1 >> 1) & 1); ++nsd1; } }
#define PNTF_WAIT_ACC0() { mbar_wait(&ctl.ar[0], nar0 & 1); ++nar0; tc_fence_after(); ++tl; trace_evt(a.trace, tron, trole, tl, 0); }
#define PNTF_WAIT_ACC1() { mbar_wait(&ctl.ar[1], nar1 & 1); ++nar1; tc_fence_after(); trace_evt(a.trace, tron, trole, tl, 2); }
#define PNTF_PUB(h) { epi_publish(&ctl.xr[h]); trace_evt(a.trace, tron, trole, tl < 0 ? 15 : tl, 1 + 2 * (h)); }
    // Bias of the epilogue that follows (loaded one epilogue earlier) and the load for the one after it.  The barrier
    // waits are volatile asm with a memory clobber, so a load written after them cannot be hoisted: issued here, before
    // the waits, its L2 latency (a tile reads 10 KB of biases, more than L1 keeps next to 213 KB of shared memory)
    // overlaps the wait and the epilogue instead of preceding every one of the 20 epilogues of a tile.
#define PNTF_BIAS(next_ptr) { bcur = bnext; bnext = __ldg(next_ptr); }
    float bcur = 0.f, bnext = __ldg(prm + c_b[0] + p.lane);
    const int trole = (tid == 0) ? 1 : 2;
    // One pass of a tile through the network: Fourier features .. output layer; leaves the output layer's
    // pre-activation rows in y13s[par] (visible to all epilogue threads on return).
    auto pass = [&](bool first, int par, bool tron_tile) {
      int tl = -1;
      const bool tron = (tid == 0 || tid == kEpiThreads - 1) && tron_tile;
      if (!first) {   // buffer 0 was last occupied by X12 of the previous tile (sd[1] also delivers X11's second half)
        PNTF_WAIT_SD0();
        PNTF_WAIT_SD1();
        PNTF_WAIT_SD1();
      }
      if constexpr (COORD_TMA) { mbar_wait(&ctl.cbar[par], (ntile >> 1) & 1); ++ntile; }   // this tile's coordinates have landed
      fourier_tile<C>(a, ts[par], buf0);
      PNTF_PUB(0);
      PNTF_PUB(1);
      // encoder: L0 ; {L1, L2(+skip)} ; {L3, L4(+skip)} ; L5 + combine
      PNTF_BIAS(prm + c_b[1] + p.lane);
      PNTF_WAIT_ACC0();
      if (!first) PNTF_WAIT_SD0();    // X13 of the previous tile
      epi_hidden_half<C, true, true>(p, ACC_A, ACC_B, p.lane, bcur, buf1, compat);
      PNTF_PUB(0);
      PNTF_BIAS(prm + c_b[2] + p.lane);
      PNTF_WAIT_ACC0();
      PNTF_WAIT_SD0();
      epi_hidden_half<C, true, false>(p, ACC_A, 0, p.lane, bcur, buf0, false);
      PNTF_PUB(0);
      PNTF_BIAS(prm + c_b[3] + p.lane);
      PNTF_WAIT_ACC0();
      PNTF_WAIT_SD0();
      epi_hidden_half<C, true, true>(p, ACC_B, ACC_B, p.lane, bcur, buf1, false);
      PNTF_PUB(0);
      PNTF_BIAS(prm + c_b[4] + p.lane);
      PNTF_WAIT_ACC0();
      PNTF_WAIT_SD0();
      epi_hidden_half<C, true, false>(p, ACC_A, 0, p.lane, bcur, buf0, false);
      PNTF_PUB(0);
      PNTF_BIAS(prm + c_b[5] + p.lane);
      PNTF_WAIT_ACC0();
      PNTF_WAIT_SD0();
      epi_hidden_half<C, true, false>(p, ACC_B, 0, p.lane, bcur, buf1, false);
      PNTF_PUB(0);
      PNTF_BIAS(prm + c_b[6] + p.lane);
      PNTF_WAIT_ACC0();
      PNTF_WAIT_SD0();
      PNTF_WAIT_SD1();                // X0's second half (buffer 0)
      epi_combine<C>(p, ACC_A, ACC_B, bcur, buf0, TRAIN ? buf1 + kHalfTileBytes : 0u);
      PNTF_PUB(0);
      PNTF_PUB(1);
      // generator: 3 x {La, Lb(+skip)} ; L12 ; L13
#pragma unroll 1
      for (int ii = 0; ii < 3; ++ii) {
        const float* ba = prm + c_b[6 + 2 * ii] + p.lane;
        const float* bb = prm + c_b[7 + 2 * ii] + p.lane;
        PNTF_BIAS(ba + 128);
        PNTF_WAIT_ACC0();
        PNTF_WAIT_SD0();
        epi_hidden_half<C, false, false>(p, ACC_A, 0, p.lane, bcur, buf1, false);
        PNTF_PUB(0);
        PNTF_BIAS(bb);
        PNTF_WAIT_ACC1();
        PNTF_WAIT_SD1();              // ii = 0: the Y5 staging image
        epi_hidden_half<C, false, false>(p, ACC_A + 128, 0, 128 + p.lane, bcur, buf1, false);
        PNTF_PUB(1);
        if (ii < 2) {
          PNTF_BIAS(bb + 128);
          PNTF_WAIT_ACC0();
          PNTF_WAIT_SD0();
          epi_hidden_half<C, false, true>(p, ACC_B, ACC_B, p.lane, bcur, buf0, false);
          PNTF_PUB(0);
          PNTF_BIAS(prm + c_b[8 + 2 * ii] + p.lane);
          PNTF_WAIT_ACC1();
          PNTF_WAIT_SD1();
          epi_hidden_half<C, false, true>(p, ACC_B + 128, ACC_B + 128, 128 + p.lane, bcur, buf0, false);
          PNTF_PUB(1);
        } else {
          PNTF_BIAS(bb + 128);
          PNTF_WAIT_ACC0();
          PNTF_WAIT_SD0();
          epi_hidden_half<C, false, false>(p, ACC_B, 0, p.lane, bcur, buf0, false);
          PNTF_PUB(0);
          PNTF_BIAS(prm + c_b[12] + p.lane);
          PNTF_WAIT_ACC1();
          PNTF_WAIT_SD1();
          epi_hidden_half<C, false, false>(p, ACC_B + 128, 0, 128 + p.lane, bcur, buf0, false);
          PNTF_PUB(1);
        }
      }
      PNTF_BIAS(prm + c_b[0] + p.lane);          // the next tile's first layer
      PNTF_WAIT_ACC0();
      PNTF_WAIT_SD0();
      epi_hidden_half<C, false, false>(p, ACC_A, 0, p.lane, bcur, buf1, false);
      PNTF_PUB(0);
      PNTF_WAIT_ACC0();
      // output layer: lane 0 of the accumulator holds y13 for all rows
      if ((warp & 3) == 0) {
        float v[32];
        tmem_ld<32>(p.tmem + ACC_A + 32 * p.rq, v);
        if (lane_id == 0) {
#pragma unroll
          for (int c = 0; c < 32; ++c) y13s[par][32 * p.rq + c] = v[c];
        }
      }
      tc_fence_before();
      named_bar_sync(1, kEpiThreads);
    };
    if constexpr (PLAN) {
      // ---- batched bidirectional path integration: per tile of P queries iterate XP += eta * Gradient(XP) until the
      // endpoints of every query met (|x1 - x0| <= thr) or max_iter + 1 steps were taken (test/gib_plan.py:74-86)
      // A tile is P query slots; a slot whose query finished takes the next unstarted query from a global counter before
      // the next pass, so every pass carries P live queries until the queue runs dry (step counts differ widely between
      // queries; rows of a tile are independent, so a query's waypoints do not depend on its tile mates).
      struct Done { int64_t q; int steps; float x[2 * D]; };
      __shared__ int active[P];
      __shared__ int steps_s[P];
      __shared__ int64_t qid[P];
      __shared__ Done done[P];
      __shared__ int ndone[2];
      const int64_t wstride = int64_t(a.max_iter + 2) * 2 * D;
      const int nrow = a.max_iter + 2;
      TcTile<C>& t0 = ts[0];
      if (tid < P) active[tid] = 0;
      if (tid == 0) { ndone[0] = 0; ndone[1] = 0; }
      bool drained = false;   // the query counter ran past the end: this slot stays empty
#pragma unroll 1
      for (uint32_t round = 0;; ++round) {
        if (tid < P && !active[tid] && !drained) {
          // refill this slot
          for (;;) {
            const int64_t q = int64_t(atomicAdd(a.tile_counter, 1));
            if (q >= a.pair_count) {
              drained = true;
              t0.valid[tid] = 0;
              t0.env[tid] = 0;
#pragma unroll
              for (int i = 0; i < 2 * D; ++i) t0.x[tid][i] = 0.f;
              break;
            }
            const int64_t pair = a.pair_begin + (a.order != nullptr ? int64_t(__ldg(a.order + q)) : q);
            float* w = a.waypoints + pair * wstride;
            float d2 = 0.f;
#pragma unroll
            for (int i = 0; i < 2 * D; ++i) {
              t0.x[tid][i] = __ldg(a.coords + src_row(a, pair) * a.cstride + i);
              w[i] = t0.x[tid][i];
            }
#pragma unroll
            for (int i = 0; i < D; ++i) { const float dv = t0.x[tid][D + i] - t0.x[tid][i]; d2 = fmaf(dv, dv, d2); }
            if (sqrtf(d2) > a.thr) {
              t0.valid[tid] = 1;
              t0.env[tid] = int(pair / a.n_per_env);
              qid[tid] = pair;
              steps_s[tid] = 0;
              active[tid] = 1;
              break;
            }
            // the endpoints already meet: no step; every row repeats the query
            if (a.fill_tail)
              for (int r = 1; r < nrow; ++r)
                for (int i = 0; i < 2 * D; ++i) w[int64_t(r) * 2 * D + i] = t0.x[tid][i];
            a.nsteps[pair] = 0;
          }
        }
        named_bar_sync(1, kEpiThreads);
        int any = 0;
#pragma unroll
        for (int q = 0; q < P; ++q) any |= active[q];
        if (!any) break;
        if (tid == 0) {
          ctl.more = 1;
          mbar_arrive(&ctl.go);
        }
        pass(false, 0, false);
        if (tid < P && active[tid]) {
          float y[CGP];
#pragma unroll
          for (int c = 0; c < CGP; ++c) y[c] = y13s[0][C::pair_row(tid) + c];
          y[0] += __ldg(prm + c_b[13]);
          const ActOut o = actout_eval(y[0]);
          float dtau[C::N1], dir[2 * D];
#pragma unroll
          for (int i = 0; i < C::N1; ++i) dtau[i] = y[1 + i] * o.o1;
          field_direction<D>(t0.x[tid], o.o, dtau, dir);
          float d2 = 0.f;
#pragma unroll
          for (int i = 0; i < 2 * D; ++i) t0.x[tid][i] = fmaf(a.eta, dir[i], t0.x[tid][i]);
#pragma unroll
          for (int i = 0; i < D; ++i) { const float dv = t0.x[tid][D + i] - t0.x[tid][i]; d2 = fmaf(dv, dv, d2); }
          const int it = ++steps_s[tid];
          float* w = a.waypoints + qid[tid] * wstride + int64_t(it) * 2 * D;
          for (int i = 0; i < 2 * D; ++i) w[i] = t0.x[tid][i];
          if (!(sqrtf(d2) > a.thr) || it == a.max_iter + 1) {
            active[tid] = 0;
            Done& r = done[atomicAdd(&ndone[round & 1], 1)];
            r.q = qid[tid];
            r.steps = it;
#pragma unroll
            for (int i = 0; i < 2 * D; ++i) r.x[i] = t0.x[tid][i];
          }
        }
        named_bar_sync(1, kEpiThreads);
        // rows after the last step of the queries that finished in this round repeat the final position (all threads);
        // the other counter was last read in the previous round, before this round's pass
        const int nd = ndone[round & 1];
        if (tid == 0) ndone[(round + 1) & 1] = 0;
        for (int k = 0; k < nd; ++k) {
          const Done& r = done[k];
          float* w = a.waypoints + r.q * wstride;
          const int first_e = (r.steps + 1) * 2 * D, last_e = a.fill_tail ? nrow * 2 * D : 0;
          for (int e = first_e + tid; e < last_e; e += kEpiThreads) __stcs(w + e, r.x[e % (2 * D)]);   // streaming: never re-read here
          if (tid == 0) a.nsteps[r.q] = r.steps;
        }
      }
      if (tid == 0) {
        ctl.more = 0;
        mbar_arrive(&ctl.go);
      }
    } else {
    // TC_TAIL_LOSS: the observed speeds of a tile reach shared memory one tile ahead of the tail that uses them (and the
    // gather index, if any, two tiles ahead), by asynchronous copies that occupy no registers: a global round trip in the
    // tail delays warp 0's share of the next tile's Fourier features and with it the whole next tile.  Neutral for the
    // 1 Mi-pair step, which is limited by the board's power budget (profiles/r02v_tail_loads.txt); kept for the cases that
    // leave power to spare, where the forward is bound by this chain.
    constexpr bool SPEED_AHEAD = (TAIL == TC_TAIL_LOSS && ORDER >= 2);
    __shared__ float spd_s[2][P][2];
    __shared__ int32_t row_s[2][P];
    auto pair_of = [&](int64_t tile) -> int64_t {     // this thread's pair in `tile`, or -1
      const int64_t pr = a.pair_begin + tile * P + tid;
      return (tile < ntiles && pr < a.pair_begin + a.pair_count && pr < a.n_total) ? pr : -1;
    };
    auto fetch_speed = [&](int64_t pr, int64_t row, int buf) {
      if (pr < 0) return;
      cp_async4(&spd_s[buf][tid][0], a.speed + row * a.sstride);
      cp_async4(&spd_s[buf][tid][1], a.speed + row * a.sstride + 1);
    };
    if constexpr (SPEED_AHEAD) {
      if (tid < P) {
        const int64_t p0 = pair_of(blockIdx.x), p1 = pair_of(int64_t(blockIdx.x) + gridDim.x);
        fetch_speed(p0, p0 >= 0 ? src_row(a, p0) : 0, 0);
        if (a.index != nullptr && p1 >= 0) cp_async4(&row_s[1][tid], a.index + p1);
      }
    }
    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x, par ^= 1) {
      const int64_t pair0 = a.pair_begin + tile * P;
      pass(tile == int64_t(blockIdx.x), par, blockIdx.x == 0 && tile == int64_t(blockIdx.x) + 2 * gridDim.x);
      if constexpr (TRAIN) {
        if (tid >= 128 && tid < 128 + kRows) a.st.y13[size_t(tile) * kRows + (tid - 128)] = y13s[par][tid - 128];
      }
      if (tid < P) {
        if constexpr (SPEED_AHEAD) cp_async_wait_all();   // this tile's speeds and the next tile's gather index (issued a tile ago)
        if (ts[par].valid[tid]) {
          const int64_t pair = pair0 + tid;
          if constexpr (ORDER == 0) {
            // value-only forward: tau, and |x1 - x0| / tau for Model.TravelTimes (S:1169-1188)
            const float tau = actout_eval(y13s[par][C::pair_row(tid)] + __ldg(prm + c_b[13])).o;
            if (a.tau) a.tau[pair] = tau;
            if constexpr (TAIL == TC_TAIL_FIELD) a.field[pair] = field_travel_time<D>(ts[par].x[tid], tau);
          } else {
          float y[C::CG];
#pragma unroll
          for (int c = 0; c < C::CG; ++c) y[c] = y13s[par][C::pair_row(tid) + c];
          y[0] += __ldg(prm + c_b[13]);
          const ActOut o = actout_eval(y[0]);
          const float tau = o.o;
          float dtau[C::N1], ltau[C::N1];
#pragma unroll
          for (int i = 0; i < C::N1; ++i) {
            const float dy = y[1 + i];
            dtau[i] = dy * o.o1;
            ltau[i] = 0.f;
            if constexpr (ORDER == 2) ltau[i] = fmaf(dy * dy, o.o2, y[1 + C::N1 + i] * kLapUnscale * o.o1);
          }
          if constexpr (C::LS) {
            // summed Laplacian columns: the per-endpoint sums go to the first slot of each endpoint (the residual
            // only uses lap_k = sum_i ltau[k D + i])
#pragma unroll
            for (int k = 0; k < 2; ++k) {
              float q = 0.f;
#pragma unroll
              for (int i = 0; i < D; ++i) q = fmaf(y[1 + k * D + i], y[1 + k * D + i], q);
              ltau[k * D] = fmaf(q, o.o2, y[1 + C::N1 + k] * kLapUnscale * o.o1);
            }
          }
          if (a.tau) a.tau[pair] = tau;
          if (a.dtau)
            for (int i = 0; i < C::N1; ++i) a.dtau[pair * C::N1 + i] = dtau[i];
          if constexpr (ORDER == 2) {
            if (a.ltau)
              for (int i = 0; i < C::N1; ++i) a.ltau[pair * C::N1 + i] = ltau[i];
          }
          const float* x = ts[par].x[tid];
          if constexpr (TAIL == TC_TAIL_LOSS && ORDER >= 2) {
            const float al = sched_alpha(a);
            const float y0 = al * spd_s[par][tid][0] + 1.0f - al;      // landed: cp_async_wait_all() above
            const float y1 = al * spd_s[par][tid][1] + 1.0f - al;
            a.diff[pair] = eikonal_residual<D, false>(x, tau, dtau, ltau, y0, y1, sched_gamma(a), a.single_variant, 0.f, nullptr,
                                                      nullptr, nullptr);
          }
          if constexpr (TAIL == TC_TAIL_FIELD) {
            if (a.field_kind == 0) a.field[pair] = field_travel_time<D>(x, tau);
            if (a.field_kind == 1) a.field[pair] = field_speed<D>(x, tau, dtau, ltau, 0.f, false);
            if (a.field_kind == 3) {
              float out[2 * D];
              field_direction<D>(x, tau, dtau, out);
              for (int i = 0; i < 2 * D; ++i) a.field[pair * 2 * D + i] = out[i];
            }
            if constexpr (ORDER >= 2) {
              if (a.field_kind == 2) a.field[pair] = field_speed<D>(x, tau, dtau, ltau, a.gamma, true);
            }
          }
          }
        }
        // this buffer's next user is the tile after next: COORD_TMA -> the MMA warp refills it (load_coords); else fetch
        // its coordinates now (the barrier of the next tile orders these writes before its readers)
        if constexpr (!COORD_TMA) load_tile(tile + 2 * int64_t(gridDim.x), ts[par]);
        if constexpr (SPEED_AHEAD) {
          const int64_t pn = pair_of(tile + gridDim.x), pn2 = pair_of(tile + 2 * int64_t(gridDim.x));
          fetch_speed(pn, (a.index != nullptr && pn >= 0) ? int64_t(row_s[par ^ 1][tid]) : pn, par ^ 1);
          if (a.index != nullptr && pn2 >= 0) cp_async4(&row_s[par][tid], a.index + pn2);
        }
      }
    }
    }
#undef PNTF_WAIT_ACC0
#undef PNTF_WAIT_ACC1
#undef PNTF_WAIT_SD0
#undef PNTF_WAIT_SD1
#undef PNTF_PUB
#undef PNTF_BIAS
  }
  tc_fence_before();
  __syncthreads();
  if (warp == kMmaWarp) tmem_dealloc(tmem_base, 512);
}

}  // namespace tc
}  // namespace pntf
