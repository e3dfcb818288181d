// The DDP Reducer + GradScaler + SGD tail of the step, as ONE launch of one 8-CTA cluster:
//   * finish the weight-gradient partial sums, round to the autocast dtype (autocast keeps
//     conv/linear weight gradients in that dtype: autocast_mode.h:824,840)
//   * Reducer: bucket = grad * (1/W), all-reduce SUM        reducer.hpp:276, ddp_main.py:121
//       -> one-shot push over NVLink peer memory into every rank's inbox as self-validating
//          {value, epoch} words (no fence, no flags), summed in fixed rank order so every replica
//          gets identical bits
//   * GradScaler.unscale_: grad *= float(1/double(scale)), found_inf    grad_scaler.py:351-356
//   * GradScaler.step: SGD p += (-lr) * g unless found_inf             grad_scaler.py:371, sgd.py:473
//   * GradScaler.update: `_amp_update_scale_`                           grad_scaler.py:549-556
//   * DDP's rank-0 broadcast of the BN running buffers (plain-BN DDP)   distributed.py:2243-2257
//   * loss = mean(nll) in the loss dtype                                ddp_main.py:89
// 8 x 1024 threads x 4 floats cover the 29,034-element bucket, so every gradient element
// stays in registers from the partial sums to the parameter update; the only grid-wide
// dependency (found_inf) is resolved through distributed shared memory.
#pragma once
#include <cooperative_groups.h>

#include "fwd_kernels.cuh"

namespace dstep {
namespace cg = cooperative_groups;

struct StepArgs {
  float* params;
  float* grads;         // tap T-c
  float* grads_scaled;  // tap T-b
  const float* gpart;   // [rows][N_PAD]
  const float* gred;    // [N_PAD] conv2-weight partial rows already summed (k_reduce_rows), or null; rows_conv2b == 0:
                        // the conv2 bias gradient as well
  const float* gdirect; // [N_PAD]
  const float* nll;     // [B]
  volatile float* host_mail;  // pinned host memory {loss, step counter as float bits}: lets the host API return the
                        // loss by polling instead of a D2H copy + stream synchronisation; may be null
  float* bnbuf;
  long long* nbt;
  DevState* ds;
  CommDev comm;
  int B;
  int rows_conv1, rows_conv2w, rows_conv2b, rows_fc;
  int use_scaler, sync;
  float growth_factor, backoff_factor;
  int growth_interval;
  // lr and the other torch.optim.SGD knobs (sgd.py:425-465) are read from DevState: run-time state, see common.cuh
  float* mom;  // [N_PAD] momentum buffers
  int post_loss;  // 1: post {loss, step counter} to host_mail (ddpstep_train_step_host polls it)
  float* acc;     // [N_PAD] accumulation buffer of DDP.no_sync() micro-steps
  int accumulate; // 1: micro-step inside no_sync(): add the local gradient to `acc`, exchange / apply nothing
  llword* tail_flags;  // [32] per-CTA flag words of the non-cluster launch shape (k_step_grid)
  int early_push; // world > 1: the conv2-filter and Linear slices were pushed by the side branches (k_reduce_rows,
                  // k_fc_reduce_push); this kernel only gathers them
};

// conv2's filter is kept, besides its fp32 master copy, as ready-made tcgen05 B operands
// (conv2_tc.cuh): 25 taps x [K chunk][N row][8] in the autocast dtype, once for the forward
// (N = co, K = ci) and once for the data gradient (N = ci, K = co, taps flipped).  They are
// rebuilt from the fp32 weights at the head of every step -- by 200 extra CTAs of the conv1
// kernel at small batch, by k_refresh_w2_tiles on a side branch of the graph when conv1 runs on
// the tensor cores -- so the conv kernels fetch their 25 KB with a single TMA copy.  (The reference re-casts every weight on every forward: autocast_mode.h:824.)
// Doing this from inside the SGD kernel was measured at +7 us: 25,600 scattered 2-byte stores
// issued by 4 SMs.
// Sum the conv2 weight-gradient partial rows (one per wgrad CTA) in fixed row order, spread
// over many SMs; runs on a side branch of the step graph, off the critical path, so that the
// 8-CTA tail kernel reads one row instead of pulling rows x 51 KB through 8 SMs' L2 ports.
// World > 1 ("early push"): this kernel and k_fc_reduce_push below also round their slice of the gradient to the
// autocast dtype, pre-divide it by W (reducer.hpp:276) and PUSH it into every rank's inbox -- 98 % of the bucket is on
// the NVLink wire 10-25 us before the tail kernel starts, which then only gathers it.  The tail keeps the 544 elements
// that exist last (conv1, the BN affine gradients, the conv2 bias).
struct PushArgs {
  const float* gpart;
  float* gred;
  int rows, rows_bias;
  CommDev comm;
  DevState* ds;
  int push;  // world > 1
  const float* acc;  // [N_PAD] no_sync accumulation buffer (added when DevState.acc_valid)
};

// one float4 of the local gradient -> tagged words in every rank's inbox (this rank's own included)
__device__ __forceinline__ void push_grad4(const CommDev& comm, const DevState* ds, const float* acc, int e, float4 g) {
  if (ds->acc_valid) {  // gradient accumulated by no_sync micro-steps (fp32, like `.grad`)
    const float4 ac = *reinterpret_cast<const float4*>(acc + e);
    g.x += ac.x; g.y += ac.y; g.z += ac.z; g.w += ac.w;
  }
  const unsigned long long epoch = ds->sync_epoch;
  const int par = (int)(epoch & 1ull);
  const unsigned int tag = (unsigned int)epoch + 1u;
  const float inv_w = 1.0f / (float)comm.world;
  const ulonglong2 w01 = make_ulonglong2(ll_pack(g.x * inv_w, tag), ll_pack(g.y * inv_w, tag));
  const ulonglong2 w23 = make_ulonglong2(ll_pack(g.z * inv_w, tag), ll_pack(g.w * inv_w, tag));
  for (int q = 0; q < comm.world; ++q) {
    ulonglong2* dst = reinterpret_cast<ulonglong2*>(&comm.peer[q]->ll_inbox[par][comm.rank][e]);
    asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" ::"l"(dst), "l"(w01.x), "l"(w01.y) : "memory");
    asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" ::"l"(dst + 1), "l"(w23.x), "l"(w23.y) : "memory");
  }
}

template <typename T>
__global__ void __launch_bounds__(128) k_reduce_rows(PushArgs a) {
  if (threadIdx.x == 0) TRACE(14 * 16 + 0);
  const int i = (blockIdx.x * 128 + threadIdx.x) * 4;
  int rows = a.rows;
  if (i >= 12800 + (a.rows_bias > 0 ? 32 : 0)) return;
  if (i >= 12800) rows = a.rows_bias;  // the conv2 bias gradient right behind the filter (bwd_mid_tc.cuh writes 2B rows)
  const float* p = a.gpart + OFF_W2 + i;
  float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int r0 = 0; r0 < rows; r0 += 16) {  // sixteen predicated loads per trip to L2, row order kept
    float4 v[16];
#pragma unroll
    for (int i = 0; i < 16; ++i)
      v[i] = r0 + i < rows ? __ldcg(reinterpret_cast<const float4*>(p + (size_t)(r0 + i) * N_PAD)) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int i = 0; i < 16; ++i) { g.x += v[i].x; g.y += v[i].y; g.z += v[i].z; g.w += v[i].w; }
  }
  *reinterpret_cast<float4*>(a.gred + OFF_W2 + i) = g;
  if (a.push && i < 12800)  // the bias tail (rows_bias) stays with the tail kernel: its region is not an early-push one
    push_grad4(a.comm, a.ds, a.acc, OFF_W2 + i, make_float4(rnd<T>(g.x), rnd<T>(g.y), rnd<T>(g.z), rnd<T>(g.w)));
  if (threadIdx.x == 0) TRACE(14 * 16 + 15);
}

// The Linear layer's weight / bias gradient (elements OFF_WFC .. N_PARAM): sum the k_fc_wgrad partial rows in row
// order, round, pre-divide, push.  World > 1 only; runs on the fc side branch of the step graph.
template <typename T>
__global__ void __launch_bounds__(128) k_fc_reduce_push(PushArgs a) {
  if (threadIdx.x == 0) TRACE(15 * 16 + 0);
  const int e = OFF_WFC + (blockIdx.x * 128 + threadIdx.x) * 4;
  if (e >= N_PAD) return;
  const float* p = a.gpart + e;
  float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int r0 = 0; r0 < a.rows; r0 += 8) {
    float4 v[8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
      v[i] = r0 + i < a.rows ? __ldcg(reinterpret_cast<const float4*>(p + (size_t)(r0 + i) * N_PAD)) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int i = 0; i < 8; ++i) { g.x += v[i].x; g.y += v[i].y; g.z += v[i].z; g.w += v[i].w; }
  }
  g = make_float4(rnd<T>(g.x), rnd<T>(g.y), rnd<T>(g.z), rnd<T>(g.w));
  if (e + 4 > N_PARAM) { g.z = 0.f; g.w = 0.f; }  // 29032..29035: two elements + two of padding
  push_grad4(a.comm, a.ds, a.acc, e, g);
  if (threadIdx.x == 0) TRACE(15 * 16 + 15);
}

template <typename T>
__global__ void __launch_bounds__(256) k_refresh_w2_tiles(const float* __restrict__ params, void* tiles) {
  w2_tile_gather<T>(params, tiles, blockIdx.x * 256 + threadIdx.x);  // one thread per tile element, coalesced stores
}

// Two launch shapes share this body:
//   CLUSTER = true : one 8-CTA cluster of 1024 threads; the step-wide flags (non-finite, peer timeout) travel through
//                    distributed shared memory and one cluster barrier.  Needs 8 completely free SMs at once.
//   CLUSTER = false: 29 ordinary CTAs of 256 threads; every CTA publishes its flags as one tagged word in global memory and
//                    polls the others' (one L2 round trip).  Small CTAs fit beside the still-running predecessor
//                    kernels, so under programmatic launch the tail is resident early instead of waiting for eight
//                    whole SMs to drain -- which is what delayed it by ~3 us at world 8, where the predecessor's CTAs
//                    spin on a cross-GPU exchange (profiles/r02_trace_b32_w8.txt).
constexpr int STEP_GRID_THREADS = 256;
constexpr int STEP_GRID_CTAS = (N_PAD + STEP_GRID_THREADS * 4 - 1) / (STEP_GRID_THREADS * 4);  // 29

template <typename T, int NTHREADS, bool CLUSTER>
__device__ __forceinline__ void step_body(const StepArgs& a) {
  if (threadIdx.x == 0) TRACE(13 * 16 + 0);
  constexpr int LANES = NTHREADS / 104;           // row lanes of the cooperative conv1-region sum (9 or 2)
  constexpr int PER = LANES >= 8 ? 4 : 16;        // loads in flight per lane and trip to L2
  __shared__ int s_nonfinite;
  __shared__ int s_flags[STEP_CTAS];
  const int tid = threadIdx.x, cta = blockIdx.x;
  const int e = (cta * NTHREADS + tid) * 4;
  const bool valid = e < N_PARAM;  // last float4 (29032..29035) is half padding
  const int W = a.comm.world, rank = a.comm.rank;
  const unsigned long long epoch = a.ds->epoch, sync_epoch = a.ds->sync_epoch;
  const int acc_valid = a.ds->acc_valid;
  const float scale = a.ds->scale;
  const int tracker = a.ds->growth_tracker;
  const int mom_ready = a.ds->mom_ready;
  // this thread's four elements belong to one parameter tensor; its param group's settings (torch/optim/sgd.py:425-465)
  const DevState::SgdGroup grp = a.ds->grp[tensor_of(min(e, N_PARAM - 1))];
  const float lr = grp.lr, momentum = grp.momentum, weight_decay = grp.weight_decay, dampening = grp.dampening;
  const int nesterov = grp.nesterov;
  // "does any group use momentum" decides whether the buffers count as initialised after this step; the thread that
  // writes the scalar state fetches the ten values NOW (independent loads, long before they are needed) -- read at the
  // end of the kernel they were ten serialised trips to L2 on the critical path (+1.9 us per step, measured)
  bool any_momentum = false;
  if (cta == 0 && tid == 0) {
    float m[10];
#pragma unroll
    for (int t = 0; t < 10; ++t) m[t] = a.ds->grp[t].momentum;
#pragma unroll
    for (int t = 0; t < 10; ++t) any_momentum = any_momentum | (m[t] != 0.f);
  }
  if (tid == 0) s_nonfinite = 0;
  // parameters are fetched now, long before they are needed, so the load overlaps everything below
  float4 p4 = make_float4(0.f, 0.f, 0.f, 0.f);
  if (valid) {
    if (e + 4 <= N_PARAM) p4 = *reinterpret_cast<const float4*>(a.params + e);
    else { p4.x = a.params[e]; p4.y = a.params[e + 1]; }
  }
  pdl_sync();  // predecessor: conv1 weight gradient (the side branches are joined by full dependencies)
  if (threadIdx.x == 0) TRACE(13 * 16 + 1);
  // a cross-rank wait of an earlier kernel of this step (or of an earlier step) timed out: bit 1 of the cluster-wide flag
  const int err_before = *reinterpret_cast<volatile int*>(&a.ds->error);
  // CTA 0 owns the conv1 region (elements 0..415), whose gradient arrives as one partial row per
  // conv1-wgrad CTA: all 1024 threads sum it together (9 row lanes per float4, one round trip to
  // L2 instead of rows/8), fixed order.  Warp 0 also fetches the per-image losses now.
  __shared__ float4 s_c1[LANES][104];
  float nll_pref = 0.f;
  // CTA 0's other elements (BN affine gradients, the head of conv2's filter) need one row each: that load is issued
  // before the cooperative sum so both travel to L2 together
  float4 pre = make_float4(0.f, 0.f, 0.f, 0.f);
  const bool pre_direct = cta == 0 && e >= OFF_G1 && e < OFF_W2;
  const bool pre_red = cta == 0 && e >= OFF_W2 && a.gred != nullptr;  // CTA 0 ends at element 4095 < OFF_B2
  if (pre_direct) pre = __ldcg(reinterpret_cast<const float4*>(a.gdirect + e));
  if (pre_red) pre = __ldcg(reinterpret_cast<const float4*>(a.gred + e));
  if (cta == 0) {
    const int e4 = tid % 104, rl = tid / 104;
    if (rl < LANES) {
      // PER predicated loads per trip, never a one-load-at-a-time remainder loop (a partially unrolled loop with
      // a run-time trip count gets one, and each of its loads is a full trip to L2)
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int r0 = rl; r0 < a.rows_conv1; r0 += PER * LANES) {
        float4 v[PER];
#pragma unroll
        for (int i = 0; i < PER; ++i)
          v[i] = r0 + LANES * i < a.rows_conv1 ? __ldcg(reinterpret_cast<const float4*>(a.gpart + (size_t)(r0 + LANES * i) * N_PAD + 4 * e4))
                                               : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int i = 0; i < PER; ++i) { acc.x += v[i].x; acc.y += v[i].y; acc.z += v[i].z; acc.w += v[i].w; }
      }
      s_c1[rl][e4] = acc;
    }
    if (tid < 32) for (int i = tid; i < a.B; i += 32) nll_pref += a.nll[i];
  }
  __syncthreads();

  // early push (world > 1): this thread's four elements are already in every inbox
  const bool pushed = a.early_push && W > 1 && ((e >= OFF_W2 && e < OFF_B2) || e >= OFF_WFC);
  // ---- 1. finish the local gradient (fixed row order => deterministic)
  float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
  if (valid && !pushed) {
    int rows;
    bool round_t = true;
    if (e < OFF_G1) rows = a.rows_conv1;
    else if (e < OFF_W2) { rows = 0; round_t = false; }
    else if (e < OFF_B2) rows = a.rows_conv2w;
    else if (e < OFF_G2) rows = a.rows_conv2b;
    else if (e < OFF_WFC) { rows = 0; round_t = false; }
    else rows = a.rows_fc;
    if (pre_direct) {
      g = pre;
    } else if (pre_red) {
      g = make_float4(rnd<T>(pre.x), rnd<T>(pre.y), rnd<T>(pre.z), rnd<T>(pre.w));
    } else if (round_t) {
      const float* p = a.gpart + e;
      if (a.gred != nullptr && e >= OFF_W2 && e < (a.rows_conv2b == 0 ? OFF_G2 : OFF_B2)) { p = a.gred + e; rows = 1; }
      if (e < OFF_G1) {  // conv1 region: lanes summed above, combine in fixed order
        rows = 0;
#pragma unroll
        for (int l = 0; l < LANES; ++l) {
          const float4 v = s_c1[l][e >> 2];
          g.x += v.x; g.y += v.y; g.z += v.z; g.w += v.w;
        }
      }
      for (int r0 = 0; r0 < rows; r0 += 4) {  // four predicated loads per trip (see above)
        float4 v[4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
          v[i] = r0 + i < rows ? __ldcg(reinterpret_cast<const float4*>(p + (size_t)(r0 + i) * N_PAD)) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int i = 0; i < 4; ++i) { g.x += v[i].x; g.y += v[i].y; g.z += v[i].z; g.w += v[i].w; }
      }
      g.x = rnd<T>(g.x); g.y = rnd<T>(g.y); g.z = rnd<T>(g.z); g.w = rnd<T>(g.w);
    } else {
      g = __ldcg(reinterpret_cast<const float4*>(a.gdirect + e));
    }
    if (e + 4 > N_PARAM) { g.z = 0.f; g.w = 0.f; }
  }

  if (threadIdx.x == 0) TRACE(13 * 16 + 2);
  // ---- 1b. gradient accumulation (DDP.no_sync): earlier micro-steps left their sum in `acc` (fp32, as `.grad` is)
  if (acc_valid && valid && !pushed) {
    float4 ac = make_float4(0.f, 0.f, 0.f, 0.f);
    if (e + 4 <= N_PARAM) ac = *reinterpret_cast<const float4*>(a.acc + e);
    else { ac.x = a.acc[e]; ac.y = a.acc[e + 1]; }
    g.x += ac.x; g.y += ac.y; g.z += ac.z; g.w += ac.w;
  }
  if (a.accumulate) {  // a micro-step: keep the sum, exchange nothing, apply nothing (uniform branch: kernel argument)
    if (valid) {
      if (e + 4 <= N_PARAM) {
        *reinterpret_cast<float4*>(a.acc + e) = g;
        *reinterpret_cast<float4*>(a.grads_scaled + e) = g;  // tap: `.grad` after a no_sync backward
      } else {
        a.acc[e] = g.x; a.acc[e + 1] = g.y;
        a.grads_scaled[e] = g.x; a.grads_scaled[e + 1] = g.y;
      }
    }
    if (cta == 0 && tid < 32) {
      const float s = warp_sum(nll_pref);
      if (tid == 0) {
        a.ds->loss = rnd<T>(s / (float)a.B);
        a.ds->acc_valid = 1;
        a.ds->epoch = epoch + 1ull;
        if (a.host_mail != nullptr && (a.post_loss || err_before != 0)) {
          a.host_mail[0] = rnd<T>(s / (float)a.B);
          if (err_before != 0) reinterpret_cast<volatile int*>(a.host_mail)[2] = 1;
          __threadfence_system();
          reinterpret_cast<volatile unsigned int*>(a.host_mail)[1] = (unsigned int)epoch + 1u;
        }
      }
    }
    return;
  }
  // ---- 2. DDP all-reduce: pre-divide, push to every peer's inbox, flag, wait, sum in rank order
  const float inv_w = 1.0f / (float)W;
  float4 h = make_float4(g.x * inv_w, g.y * inv_w, g.z * inv_w, g.w * inv_w);
  float4 tot = h;
  if (W > 1) {
    const int par = (int)(sync_epoch & 1ull);
    const unsigned int tag = (unsigned int)sync_epoch + 1u;
    if (valid && !pushed) {  // every element travels with its own tag: no fence, no flag, no block-level synchronisation
      const ulonglong2 w01 = make_ulonglong2(ll_pack(h.x, tag), ll_pack(h.y, tag));
      const ulonglong2 w23 = make_ulonglong2(ll_pack(h.z, tag), ll_pack(h.w, tag));
      for (int q = 0; q < W; ++q)
        if (q != rank) {
          ulonglong2* dst = reinterpret_cast<ulonglong2*>(&a.comm.peer[q]->ll_inbox[par][rank][e]);
          asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" ::"l"(dst), "l"(w01.x), "l"(w01.y) : "memory");
          asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" ::"l"(dst + 1), "l"(w23.x), "l"(w23.y) : "memory");
        }
    }
    if (!a.sync && cta == 0) {  // plain-BN DDP: rank 0's buffers win (distributed.py:2243-2257)
      if (rank == 0) {
        for (int q = 1; q < W; ++q) {
          if (tid < N_BNBUF) ll_store(&a.comm.peer[q]->ll_bnbuf[tid], a.bnbuf[tid], tag);
          if (tid < 4) ll_store(&a.comm.peer[q]->ll_bnbuf[N_BNBUF + tid],
                                __uint_as_float(reinterpret_cast<const unsigned int*>(a.nbt)[tid]), tag);
        }
      } else {
        float bv = 0.f, nv = 0.f;
        if (tid < N_BNBUF) bv = ll_wait(&a.comm.self->ll_bnbuf[tid], tag, a.ds);
        if (tid < 4) nv = ll_wait(&a.comm.self->ll_bnbuf[N_BNBUF + tid], tag, a.ds);
        if (*reinterpret_cast<volatile int*>(&a.ds->error) == 0) {  // never adopt words of a timed-out wait
          if (tid < N_BNBUF) a.bnbuf[tid] = bv;
          if (tid < 4) reinterpret_cast<unsigned int*>(a.nbt)[tid] = __float_as_uint(nv);
        }
      }
    }
    if (valid) {
      // this rank's own contribution is stored locally too, so one gather covers all W senders
      if (!pushed) {
        llword* mine = &a.comm.self->ll_inbox[par][rank][e];
        ll_store(mine, h.x, tag); ll_store(mine + 1, h.y, tag); ll_store(mine + 2, h.z, tag); ll_store(mine + 3, h.w, tag);
      }
      const int idx[4] = {0, 1, 2, 3};
      tot = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int half = 0; half < 2; ++half) {  // two gathers of four senders keep the 1024-thread CTA within 64 registers
        if (half * 4 < W) {
          float got[16];
          if (!ll_gather<4, 4>(&a.comm.self->ll_inbox[par][0][e], N_PAD, idx, half * 4, W, tag, got, a.ds)) atomicOr(&s_nonfinite, 2);
#pragma unroll
          for (int q = 0; q < 4; ++q)  // fixed rank order
            if (half * 4 + q < W) { tot.x += got[q * 4]; tot.y += got[q * 4 + 1]; tot.z += got[q * 4 + 2]; tot.w += got[q * 4 + 3]; }
        }
      }
    }
  }

  if (threadIdx.x == 0) TRACE(13 * 16 + 3);
  // ---- 3. unscale + non-finite check (every element is multiplied, finite or not)
  float4 u = tot;
  if (a.use_scaler) {
    const float inv_scale = (float)(1.0 / (double)scale);
    u = make_float4(tot.x * inv_scale, tot.y * inv_scale, tot.z * inv_scale, tot.w * inv_scale);
    const bool bad = !(isfinite(u.x) && isfinite(u.y) && isfinite(u.z) && isfinite(u.w));
    if (valid && bad) atomicOr(&s_nonfinite, 1);
  }
  if (err_before != 0 && tid == 0) atomicOr(&s_nonfinite, 2);
  // ---- 4. cluster-wide found_inf through distributed shared memory: every CTA writes its flag into
  // slot [its rank] of ALL eight CTAs (the cluster is co-scheduled, so remote shared memory exists
  // from the start), ONE cluster barrier, then everybody reads its own eight slots
  __syncthreads();
  int flags = 0;
  if constexpr (CLUSTER) {
    cg::cluster_group cluster = cg::this_cluster();
    if (tid < STEP_CTAS) *cluster.map_shared_rank(&s_flags[cta], tid) = s_nonfinite;
    cluster.sync();
#pragma unroll
    for (int r = 0; r < STEP_CTAS; ++r) flags |= s_flags[r];
  } else {
    // every CTA publishes {flags, step tag} as one 8-byte word; warp 0 polls all of them (the loads of a polling round
    // are in flight together), ORs them and hands the result to the CTA.  The tag (epoch + 1) is never reused.
    const unsigned int ftag = (unsigned int)epoch + 1u;
    if (tid == 0) ll_store(&a.tail_flags[cta], __int_as_float(s_nonfinite), ftag);
    if (tid < 32) {
      int f = 0;
      if (tid < (int)gridDim.x) f = __float_as_int(ll_wait(&a.tail_flags[tid], ftag, a.ds));
      f = __reduce_or_sync(0xffffffffu, (unsigned int)f);
      if (*reinterpret_cast<volatile int*>(&a.ds->error) != 0) f |= 2;  // the poll itself timed out
      if (tid == 0) s_flags[0] = f;
    }
    __syncthreads();
    flags = s_flags[0];
  }
  if (threadIdx.x == 0) TRACE(13 * 16 + 4);
  // bit 1: a peer did not answer in time somewhere in this step.  The gradient is then not the all-reduced one: the
  // step applies NOTHING (parameters, momentum, scaler state stay as they were) and tells the host (DDPSTEP_E_PEER).
  const bool dead = (flags & 2) != 0;
  const int found = (flags & 1) | (dead ? 1 : 0);

  // the gradient taps are stored only now: a release at cluster scope waits for every global
  // store issued before it, so none may be in flight while the barrier above runs
  if (threadIdx.x == 0) TRACE(13 * 16 + 5);
  if (valid) {
    if (e + 4 <= N_PARAM) {
      *reinterpret_cast<float4*>(a.grads_scaled + e) = tot;
      *reinterpret_cast<float4*>(a.grads + e) = u;
    } else {
      a.grads_scaled[e] = tot.x; a.grads_scaled[e + 1] = tot.y;
      a.grads[e] = u.x; a.grads[e + 1] = u.y;
    }
  }

  // ---- 5. SGD (skipped on overflow), same expression as _foreach_add_(p, g, alpha=-lr)
  if (valid && !found) {
    float4 p = p4;
    if (weight_decay != 0.f) {  // _foreach_add(grads, params, alpha=weight_decay): a new tensor, .grad stays as it was
      u.x = fmaf(weight_decay, p.x, u.x); u.y = fmaf(weight_decay, p.y, u.y);
      u.z = fmaf(weight_decay, p.z, u.z); u.w = fmaf(weight_decay, p.w, u.w);
    }
    if (momentum != 0.f) {
      float4 b = u;  // first executed step: buf = clone(grad)
      if (mom_ready) {  // _foreach_mul_(bufs, momentum); _foreach_add_(bufs, grads, alpha=1 - dampening)
        b = make_float4(0.f, 0.f, 0.f, 0.f);
        if (e + 4 <= N_PARAM) b = *reinterpret_cast<const float4*>(a.mom + e);
        else { b.x = a.mom[e]; b.y = a.mom[e + 1]; }
        const float k = 1.0f - dampening;
        b.x = fmaf(k, u.x, b.x * momentum); b.y = fmaf(k, u.y, b.y * momentum);
        b.z = fmaf(k, u.z, b.z * momentum); b.w = fmaf(k, u.w, b.w * momentum);
      }
      if (e + 4 <= N_PARAM) *reinterpret_cast<float4*>(a.mom + e) = b;
      else { a.mom[e] = b.x; a.mom[e + 1] = b.y; }
      if (nesterov) {  // _foreach_add_(grads, bufs, alpha=momentum)
        u.x = fmaf(momentum, b.x, u.x); u.y = fmaf(momentum, b.y, u.y);
        u.z = fmaf(momentum, b.z, u.z); u.w = fmaf(momentum, b.w, u.w);
      } else {
        u = b;
      }
    }
    p.x = fmaf(-lr, u.x, p.x); p.y = fmaf(-lr, u.y, p.y);
    p.z = fmaf(-lr, u.z, p.z); p.w = fmaf(-lr, u.w, p.w);
    if (e + 4 <= N_PARAM) {
      *reinterpret_cast<float4*>(a.params + e) = p;
    } else {
      a.params[e] = p.x;
      a.params[e + 1] = p.y;
    }
  }

  if (threadIdx.x == 0) TRACE(13 * 16 + 6);
  // ---- 6. scalar state: loss, scale, tracker, epoch
  if (cta == 0 && tid < 32) {
    float s = warp_sum(nll_pref);
    if (tid == 0) {
      a.ds->loss = rnd<T>(s / (float)a.B);
      a.ds->found_inf = found ? 1.0f : 0.0f;
      if (a.use_scaler && !dead) {
        if (found) {
          a.ds->scale = scale * a.backoff_factor;
          a.ds->growth_tracker = 0;
        } else {
          const int t = tracker + 1;
          if (t == a.growth_interval) {
            const float grown = scale * a.growth_factor;
            if (isfinite(grown)) a.ds->scale = grown;
            a.ds->growth_tracker = 0;
          } else {
            a.ds->growth_tracker = t;
          }
        }
      }
      if (!found && any_momentum) a.ds->mom_ready = 1;  // buffers exist from the first executed step of a group with momentum
      a.ds->epoch = epoch + 1ull;
      a.ds->sync_epoch = sync_epoch + 1ull;
      a.ds->acc_valid = 0;
      // The pinned-host mailbox costs a system-scope fence (~2 us at the very end of the step), so only the graphs of
      // the host entry point (which polls it instead of synchronising) post to it; a voided step always raises word 2.
      if (a.host_mail != nullptr && (a.post_loss || dead)) {
        a.host_mail[0] = rnd<T>(s / (float)a.B);
        if (dead) reinterpret_cast<volatile int*>(a.host_mail)[2] = 1;  // sticky: the host API refuses further steps
        __threadfence_system();
        reinterpret_cast<volatile unsigned int*>(a.host_mail)[1] = (unsigned int)epoch + 1u;
      }
    }
  }
  if (threadIdx.x == 0) TRACE(13 * 16 + 15);
}

template <typename T>
__global__ void __cluster_dims__(STEP_CTAS, 1, 1) __launch_bounds__(STEP_THREADS) k_step(StepArgs a) {
  step_body<T, STEP_THREADS, true>(a);
}
template <typename T>
__global__ void __launch_bounds__(STEP_GRID_THREADS) k_step_grid(StepArgs a) {
  step_body<T, STEP_GRID_THREADS, false>(a);
}

// ---------------------------------------------------------------------------------------
// The exchange on its own, for ANY fp32 bucket (alternative boundary B / 8f N3): what a DDP communication hook
// -- `register_comm_hook(state, hook)`, hook(state, bucket) -> Future[Tensor] (distributed.py:1987-2067) -- has to do
// for one GradBucket: bucket <- sum over ranks of bucket * (1/W)  (`_AllReduceBySumCommHook`, default_comm_hooks.hpp:36-51).
// One launch handles one chunk of up to N_PAD elements through the same double-buffered inbox of tagged words as the
// step's own gradient (fixed rank-order sum: identical bits on every rank); longer buckets are a sequence of launches.
struct BucketArgs {
  float* bucket;      // device pointer to this chunk
  int n;              // elements in this chunk (<= N_PAD)
  float pre_scale;    // 1/W to average like DDP's default hook, 1 for a plain SUM
  DevState* ds;
  CommDev comm;
};

__global__ void __launch_bounds__(256) k_bucket_allreduce(BucketArgs a) {
  const int e = (blockIdx.x * 256 + threadIdx.x) * 4;
  if (e >= a.n) return;
  const int W = a.comm.world, rank = a.comm.rank;
  const unsigned long long ep = a.ds->sync_epoch;
  const int par = (int)(ep & 1ull);
  const unsigned int tag = (unsigned int)ep + 1u;
  float v[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) v[i] = e + i < a.n ? a.bucket[e + i] * a.pre_scale : 0.f;
  if (W > 1) {
    const ulonglong2 w01 = make_ulonglong2(ll_pack(v[0], tag), ll_pack(v[1], tag));
    const ulonglong2 w23 = make_ulonglong2(ll_pack(v[2], tag), ll_pack(v[3], tag));
    for (int q = 0; q < W; ++q) {
      ulonglong2* dst = reinterpret_cast<ulonglong2*>(&a.comm.peer[q]->ll_inbox[par][rank][e]);
      asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" ::"l"(dst), "l"(w01.x), "l"(w01.y) : "memory");
      asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" ::"l"(dst + 1), "l"(w23.x), "l"(w23.y) : "memory");
    }
    const int idx[4] = {0, 1, 2, 3};
    float tot[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      if (half * 4 < W) {
        float got[16];
        ll_gather<4, 4>(&a.comm.self->ll_inbox[par][0][e], N_PAD, idx, half * 4, W, tag, got, a.ds);
#pragma unroll
        for (int q = 0; q < 4; ++q)  // fixed rank order
          if (half * 4 + q < W) { tot[0] += got[q * 4]; tot[1] += got[q * 4 + 1]; tot[2] += got[q * 4 + 2]; tot[3] += got[q * 4 + 3]; }
      }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) v[i] = tot[i];
  }
#pragma unroll
  for (int i = 0; i < 4; ++i)
    if (e + i < a.n) a.bucket[e + i] = v[i];
}
// stream-ordered after a chunk: the next chunk (or the next step) uses the other inbox buffer and a fresh tag
__global__ void k_bump_sync_epoch(DevState* ds) { ds->sync_epoch += 1ull; }

// dist.reduce(size, 0, SUM); dist.reduce(correct, 0, SUM)  (ddp_main.py:108-109): every
// rank pushes its two fp32 counters into rank 0's slots; rank 0 sums them in rank order.
struct CountsArgs {
  float* counts;
  DevState* ds;
  CommDev comm;
};

__global__ void k_reduce_counts(CountsArgs a) {
  const int tid = threadIdx.x, W = a.comm.world, rank = a.comm.rank;
  const unsigned long long ep = a.ds->eval_epoch;
  if (W > 1) {
    if (tid == 0) {
      // the slot is single-buffered: rank 0 must have consumed our previous push
      if (ep > 0) wait_flag(&a.comm.self->flag_counts_ack, ep, a.ds);
      a.comm.peer[0]->slot_counts[rank][0] = a.counts[0];
      a.comm.peer[0]->slot_counts[rank][1] = a.counts[1];
      __threadfence_system();
      st_release_sys(&a.comm.peer[0]->flag_counts[rank], ep + 1ull);
    }
    if (rank == 0) {
      if (tid < W) wait_flag(&a.comm.self->flag_counts[tid], ep + 1ull, a.ds);
      __syncthreads();
      if (tid < 2) {
        float s = 0.f;
        for (int r = 0; r < W; ++r) s += __ldcg(&a.comm.self->slot_counts[r][tid]);
        a.counts[tid] = s;
      }
      __syncthreads();
      if (tid < W) st_release_sys(&a.comm.peer[tid]->flag_counts_ack, ep + 1ull);
    }
  }
  __syncthreads();
  if (tid == 0) a.ds->eval_epoch = ep + 1ull;
}

}  // namespace dstep
