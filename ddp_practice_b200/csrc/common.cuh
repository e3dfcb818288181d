// Shared device-side definitions of the fused ConvNet DDP+AMP step (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace dstep {

// ---- flat parameter / gradient bucket layout (definition order, ddp_main.py:16-28) ----
constexpr int N_PARAM = 29034;
constexpr int N_PAD = 29056;  // padded row length of every bucket-shaped buffer (multiple of 32)
constexpr int OFF_W1 = 0, OFF_B1 = 400, OFF_G1 = 416, OFF_BE1 = 432;
constexpr int OFF_W2 = 448, OFF_B2 = 13248, OFF_G2 = 13280, OFF_BE2 = 13312;
constexpr int OFF_WFC = 13344, OFF_BFC = 29024;
// BN buffer block: rm1[16] rv1[16] rm2[32] rv2[32]
constexpr int BNB_RM1 = 0, BNB_RV1 = 16, BNB_RM2 = 32, BNB_RV2 = 64, N_BNBUF = 96;

// which of the ten parameter tensors element e of the flat vector belongs to (every float4 lies inside one tensor:
// all offsets are multiples of 4)
__host__ __device__ __forceinline__ int tensor_of(int e) {
  return (e >= OFF_B1) + (e >= OFF_G1) + (e >= OFF_BE1) + (e >= OFF_W2) + (e >= OFF_B2) + (e >= OFF_G2) + (e >= OFF_BE2) +
         (e >= OFF_WFC) + (e >= OFF_BFC);
}

constexpr int MAX_WORLD = 8;
constexpr int STEP_CTAS = 8;        // one cluster
constexpr int STEP_THREADS = 1024;  // 8 * 1024 * 4 = 32768 >= N_PAD
constexpr int MAX_PART_ROWS = 320;  // rows of the weight-gradient partial buffer

// ---- storage dtype helpers -------------------------------------------------------------
template <typename T> struct Cvt;
template <> struct Cvt<float> {
  static __device__ __forceinline__ float to_f(float v) { return v; }
  static __device__ __forceinline__ float from_f(float v) { return v; }
};
template <> struct Cvt<__half> {
  static __device__ __forceinline__ float to_f(__half v) { return __half2float(v); }
  static __device__ __forceinline__ __half from_f(float v) { return __float2half_rn(v); }
};
template <> struct Cvt<__nv_bfloat16> {
  static __device__ __forceinline__ float to_f(__nv_bfloat16 v) { return __bfloat162float(v); }
  static __device__ __forceinline__ __nv_bfloat16 from_f(float v) { return __float2bfloat16_rn(v); }
};
// round an fp32 value to the storage dtype and come back (the autocast rounding point)
template <typename T> __device__ __forceinline__ float rnd(float v) { return Cvt<T>::to_f(Cvt<T>::from_f(v)); }
template <typename T> __device__ __forceinline__ float ldf(const T* p) { return Cvt<T>::to_f(*p); }
template <typename T> __device__ __forceinline__ void stf(T* p, float v) { *p = Cvt<T>::from_f(v); }

// conv output epilogue: cudnn_convolution rounds, then add_(bias) rounds again (dbl = 1)
template <typename T> __device__ __forceinline__ float conv_out(float acc, float bias_r, int dbl) {
  return rnd<T>((dbl ? rnd<T>(acc) : acc) + bias_r);
}

// ---- where a batch comes from ------------------------------------------------------------
// Either fp32 images already gathered into a batch (x, ddp_main.py:86), or -- the input-pipeline
// row of SURVEY 8(f) -- the whole dataset resident as uint8 [N,784] plus this batch's sampler
// indices: the gather and ToTensor's `uint8 -> fp32, / 255` (transforms/functional.py:157) are
// fused into the load, so no float batch ever exists in memory.
struct BatchSrc {
  const float* x;            // [B,784] fp32, or null
  const unsigned char* u8;   // [N,784] uint8 dataset (used when x is null)
  const long long* idx;      // [B] rows of the dataset (also indexes the labels), or null
};
// four consecutive pixels (i-th float4 of image n) as fp32
__device__ __forceinline__ float4 load_pixels4(const BatchSrc& b, int n, int i) {
  if (b.x != nullptr) return __ldg(reinterpret_cast<const float4*>(b.x + (size_t)n * 784) + i);
  const uchar4 q = __ldg(reinterpret_cast<const uchar4*>(b.u8 + (size_t)b.idx[n] * 784) + i);
  return make_float4((float)q.x / 255.0f, (float)q.y / 255.0f, (float)q.z / 255.0f, (float)q.w / 255.0f);
}

// ---- cross-GPU symmetric region (identical layout on every rank) -----------------------
// Exchanges are PUSH and fence-free: every value travels as ONE naturally aligned 8-byte word
// {fp32 payload, 32-bit epoch tag} stored straight into the consumer's memory over NVLink (an
// aligned 8-byte store is single-copy atomic), and the consumer polls its OWN memory until the tag
// of the word it needs equals the epoch it is in -- the data is its own flag (the "LL" idea).  No
// __threadfence_system, no separate flag store, no ordering between words: one NVLink store latency
// per exchange.  Tags are epoch + 1 (never 0, never reused), so stale words can never match.
typedef unsigned long long llword;
constexpr int BN_FWD_SLOT = 104;  // 3 x 32 channels + count, padded
struct alignas(128) SymRegion {
  alignas(128) llword ll_bn_fwd[2][MAX_WORLD][BN_FWD_SLOT];  // [layer][sender]: mean[C] invstd[C] unbiased var[C] count
  alignas(128) llword ll_bn_bwd[2][MAX_WORLD][64];   // [layer][sender]: sum_dy[C] sum_dy_xmu[C]
  alignas(128) llword ll_bnbuf[N_BNBUF + 4];         // rank 0's running buffers + 2 x nbt (lo, hi) (plain-BN DDP)
  alignas(128) llword ll_inbox[2][MAX_WORLD][N_PAD]; // [step parity][sender][bucket element]
  // test() counters: flag + slot protocol (off the training path)
  unsigned long long flag_counts[MAX_WORLD];
  unsigned long long flag_counts_ack;                // rank 0 consumed this rank's counters
  alignas(128) float slot_counts[MAX_WORLD][2];
};

struct CommDev {
  int rank, world;
  SymRegion* self;
  SymRegion* peer[MAX_WORLD];  // peer[rank] == self
};

// device-resident step state
struct DevState {
  unsigned long long epoch;       // completed training steps (flag value base)
  unsigned long long eval_epoch;  // completed count reductions
  unsigned int ticket[8];         // "last CTA finishes" counters, one per producing kernel
  int error;                      // 0 ok, 1 peer timeout
  float scale;
  int growth_tracker;
  float found_inf;
  float loss;
  int mom_ready;                  // SGD momentum buffers hold values (set by the first executed step)
  // Run-time hyper-parameters of torch.optim.SGD (param_groups[0]; sgd.py:425-465).  They live in device memory, not in
  // kernel arguments, so an LR scheduler or a hand-edited param group changes them without a new context or graph
  // (ddpstep_set_sgd).
  // One entry per parameter tensor in definition order (conv1.w, conv1.b, bn1.g, bn1.b, conv2.w, conv2.b, bn2.g, bn2.b,
  // fc.w, fc.b), so that optimizer.param_groups with different settings are expressible (ddpstep_set_sgd_group);
  // ddpstep_set_sgd writes all ten.
  struct SgdGroup { float lr, momentum, weight_decay, dampening; int nesterov; } grp[10];
  // Cross-rank waits give up after this many SM clock cycles (ddpstep_set_peer_timeout_ms) and raise `error`; a step
  // that saw the error applies no update (k_step) and the host API reports DDPSTEP_E_PEER.
  long long timeout_cycles;
  // DDP.no_sync() / gradient accumulation (distributed.py:1441-1470): a micro-step adds its local (scaled, rounded)
  // gradient to the accumulation buffer instead of exchanging it; the next ordinary step exchanges buffer + its own.
  int acc_valid;                  // the accumulation buffer holds a gradient
  unsigned long long sync_epoch;  // completed SYNCHRONISING steps: parity / tag of the gradient inbox (micro-steps do not
                                  // touch the inbox, so `epoch` would let a rank run two buffers ahead of a peer)
};

// global BN statistics of the current step, written by the bn_relu_pool kernel
struct BnStat {
  float mean[32];
  float invstd[32];
  float n_total;
  // local (this rank) statistics from the conv epilogue
  float lmean[32];
  float linvstd[32];
  float lvar_unbiased[32];
  // backward sums, local and global
  float lsum_dy[32], lsum_dy_xmu[32];
  float sum_dy[32], sum_dy_xmu[32];
};

__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ llword ll_pack(float v, unsigned int tag) {
  return ((llword)tag << 32) | (llword)__float_as_uint(v);
}
__device__ __forceinline__ void ll_store(llword* p, float v, unsigned int tag) {
  asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(p), "l"(ll_pack(v, tag)) : "memory");
}
// Partial-sum rows between a producer kernel and the BN kernel after it come in two formats: plain fp32 (the
// consumer reads them after the grid dependency resolved) or tagged words like the cross-GPU ones ("ll rows"), which
// ONE CTA of the consumer can poll while the producer grid is still draining -- see bn_forward_publish.
__device__ __forceinline__ void store_partial(float* part, size_t i, float v, int ll, unsigned int tag) {
  if (ll) ll_store(reinterpret_cast<llword*>(part) + i, v, tag);
  else part[i] = v;
}
// A cross-rank wait that has spun for longer than the configured timeout raises the sticky error word: the step
// then applies no update (k_step) and the host API returns DDPSTEP_E_PEER -- a late or dead peer must never end in a
// silently different replica.
__device__ __forceinline__ bool spin_expired(long long t0, DevState* ds) {
  if (*reinterpret_cast<volatile int*>(&ds->error) != 0) return true;  // the step is void already: do not wait again
  if (clock64() - t0 > *reinterpret_cast<volatile long long*>(&ds->timeout_cycles)) { atomicExch(&ds->error, 1); return true; }
  return false;
}
// Poll a word of this rank's own region until its tag matches (bounded by the peer timeout).
__device__ __forceinline__ float ll_wait(const llword* p, unsigned int tag, DevState* ds) {
  llword w;
  long long t0 = 0;
  for (int spin = 0;; ++spin) {
    asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(w) : "l"(p) : "memory");
    if ((unsigned int)(w >> 32) == tag) break;
    if (spin == 64) t0 = clock64();
    if (spin > 64 && spin_expired(t0, ds)) break;
  }
  return __uint_as_float((unsigned int)w);
}

__device__ __forceinline__ llword ll_ld(const llword* p) {
  llword w;
  asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(w) : "l"(p) : "memory");
  return w;
}
// Gather NW words from each of `world` senders (sender r's word j at base + r*stride + idx[j]).  All
// loads of one polling round are issued together, so a round costs ONE L2 round trip instead of
// world x NW dependent ones; rounds repeat until every tag matches.  out[r*NW + j] = payload.
template <int NW, int NR = MAX_WORLD>
__device__ __forceinline__ bool ll_gather(const llword* base, size_t stride, const int (&idx)[NW], int r0, int world,
                                          unsigned int tag, float* out, DevState* ds) {
  long long t0 = 0;
  for (int spin = 0;; ++spin) {
    llword w[NR * NW];
#pragma unroll
    for (int r = 0; r < NR; ++r)
#pragma unroll
      for (int j = 0; j < NW; ++j)
        w[r * NW + j] = (r0 + r) < world ? ll_ld(base + (size_t)(r0 + r) * stride + idx[j]) : ((llword)tag << 32);
    bool ok = true;
#pragma unroll
    for (int i = 0; i < NR * NW; ++i) ok = ok && ((unsigned int)(w[i] >> 32) == tag);
    if (ok) {
#pragma unroll
      for (int i = 0; i < NR * NW; ++i) out[i] = __uint_as_float((unsigned int)w[i]);
      return true;
    }
    if (spin == 64) t0 = clock64();
    if (spin > 64 && spin_expired(t0, ds)) {  // hand back what is there; the error word voids the step
#pragma unroll
      for (int i = 0; i < NR * NW; ++i) out[i] = __uint_as_float((unsigned int)w[i]);
      return false;
    }
  }
}

// Spin until *flag >= want; gives up after the peer timeout and raises the error word so a dead peer
// cannot hang the GPU forever.
__device__ __forceinline__ void wait_flag(const unsigned long long* flag, unsigned long long want, DevState* ds) {
  long long t0 = clock64();
  while (ld_acquire_sys(flag) < want)
    if (spin_expired(t0, ds)) break;
}

// Programmatic dependent launch (PDL): inside the step graph a kernel may be launched while its
// predecessor is still running.  Everything before pdl_wait() may only touch memory that no kernel of
// the same step has written before it two or more launches back (parameters, BN statistics of
// earlier kernels, its own shared memory / TMEM); pdl_wait() returns when the predecessor grid has
// completed and its writes are visible.  pdl_trigger() (issued right after, so at most two grids ever
// overlap) lets the successor start its own prologue.  Both are no-ops for ordinary launches.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_sync() { pdl_wait(); pdl_trigger(); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// true in exactly one CTA of the grid: the last one to get here.  Resets the ticket so
// the same counter works on the next launch / graph replay.
__device__ __forceinline__ bool last_cta_done(unsigned int* ticket, unsigned int n_ctas) {
  __shared__ int s_last;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0 && threadIdx.y == 0) {
    unsigned int t = atomicAdd(ticket, 1u);
    s_last = (t == n_ctas - 1);
    if (s_last) *ticket = 0;
  }
  __syncthreads();
  if (s_last) __threadfence();
  return s_last != 0;
}

// ReLU(max_pool 2x2) of four BN outputs as ATen computes it: a NaN in the window wins the pool
// (`val > maxval || isnan(val)`) and passes the ReLU (`x <= 0 ? 0 : x`), so a poisoned forward reaches the loss and
// GradScaler sees it.  fmaxf alone would drop the NaN.  The sum is NaN only for a NaN or inf - inf: one compare on
// the hot path, the exact test only behind it.
__device__ __forceinline__ float relu_max4(float v0, float v1, float v2, float v3) {
  const float m = fmaxf(fmaxf(fmaxf(v0, v1), fmaxf(v2, v3)), 0.f);
  const float s = (v0 + v1) + (v2 + v3);
  if (s != s && (v0 != v0 || v1 != v1 || v2 != v2 || v3 != v3)) return __int_as_float(0x7fc00000);
  return m;
}

// The four BN outputs of one 2x2 pooling window and who wins the pool.
// Winner = first maximum of the ReLU outputs in scan order (max_pool2d_with_indices);
// gradient flows only if the winner's ReLU output is > 0 (threshold_backward).
template <typename T>
__device__ __forceinline__ int pool_winner(const float y[4], float mu, float is, float g, float b, bool& live) {
  float best = -1.f;
  int arg = 0;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float r = fmaxf(rnd<T>((y[i] - mu) * is * g + b), 0.f);
    if (r > best) { best = r; arg = i; }
  }
  live = best > 0.f;
  return arg;
}
// Same, also handing back the winner's conv output (tracked with selects: indexing y[] by the run-time winner would
// push the array into local memory, a chain of ?: over it compiles to branches).
template <typename T>
__device__ __forceinline__ int pool_winner(const float y[4], float mu, float is, float g, float b, bool& live, float& win) {
  float best = -1.f;
  int arg = 0;
  win = y[0];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float r = fmaxf(rnd<T>((y[i] - mu) * is * g + b), 0.f);
    const bool better = r > best;
    best = better ? r : best;
    arg = better ? i : arg;
    win = better ? y[i] : win;
  }
  live = best > 0.f;
  return arg;
}

// Optional in-kernel phase timestamps (build with -DDDPSTEP_TRACE): CTA 0 writes globaltimer values
// into a small global buffer so latency chains inside a kernel can be read off directly.
#ifdef DDPSTEP_TRACE
__device__ unsigned long long g_trace[256];  // kernel k owns slots 16k .. 16k+15 (0 = entry)
__device__ __forceinline__ void trace_mark(int slot) {
  if (blockIdx.x == 0 && blockIdx.y == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)::"memory");  // the clobber keeps loads / stores on their side of the mark
    g_trace[slot] = t;
  }
}
#define TRACE(slot) trace_mark(slot)
#else
#define TRACE(slot)
#endif

}  // namespace dstep
