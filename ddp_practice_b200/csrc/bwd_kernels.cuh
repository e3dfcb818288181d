// Backward kernels of the ConvNet step (direct, CUDA-core path; any batch size).
// Autograd of ddp_main.py:91 `scaler.scale(loss).backward()` through ddp_main.py:16-28.
// ReLU masks and max-pool winners are recomputed from the saved conv outputs and the BN
// statistics instead of being stored (the reference keeps int64 pool indices, 4x the data).
#pragma once
#include "fwd_kernels.cuh"

namespace dstep {

// ---------------------------------------------------------------------------------------
// fc weight / bias gradient: dW[j][k] = sum_n dlogits[n][j] * p2[n][k], db[j] = sum_n dlogits[n][j]
// grid (7 k-chunks of 224, NS batch splits); partial rows go to gpart[split].
struct FcWgradArgs {
  const void* p2;
  const void* dlogits;
  float* gpart;  // [rows][N_PAD]
  int B, per_split;
};

template <typename T>
__global__ void __launch_bounds__(224) k_fc_wgrad(FcWgradArgs a) {
  if (threadIdx.x == 0) TRACE(6 * 16 + 0);
  __shared__ float dl_s[32][10];
  const int tid = threadIdx.x;
  const int k = blockIdx.x * 224 + tid;
  const int n0 = blockIdx.y * a.per_split;
  const int n1 = min(a.B, n0 + a.per_split);
  const T* __restrict__ p2 = reinterpret_cast<const T*>(a.p2);
  const T* __restrict__ dl = reinterpret_cast<const T*>(a.dlogits);
  float acc[10], bsum = 0.f;
#pragma unroll
  for (int j = 0; j < 10; ++j) acc[j] = 0.f;
  for (int nb = n0; nb < n1; nb += 32) {
    const int cnt = min(32, n1 - nb);
    __syncthreads();
    for (int i = tid; i < cnt * 10; i += 224) dl_s[i / 10][i % 10] = ldf<T>(&dl[(size_t)nb * 10 + i]);
    __syncthreads();
    for (int i = 0; i < cnt; ++i) {
      const float x = ldf<T>(&p2[(size_t)(nb + i) * 1568 + k]);
#pragma unroll
      for (int j = 0; j < 10; ++j) acc[j] = fmaf(dl_s[i][j], x, acc[j]);
    }
    if (blockIdx.x == 0 && tid < 10)
      for (int i = 0; i < cnt; ++i) bsum += dl_s[i][tid];
  }
  float* row = a.gpart + (size_t)blockIdx.y * N_PAD;
#pragma unroll
  for (int j = 0; j < 10; ++j) row[OFF_WFC + j * 1568 + k] = acc[j];
  if (blockIdx.x == 0 && tid < 10) row[OFF_BFC + tid] = bsum;
  if (threadIdx.x == 0) TRACE(6 * 16 + 15);
}

// ---------------------------------------------------------------------------------------
// BN backward, pass 1 (`batch_norm_backward_reduce`, _functions.py:144): per channel
// sum(dy) and sum(dy * (x - mean)) where dy is the gradient arriving at the BN output, i.e.
// d(pool out) routed through max-pool and ReLU.  grid (C, NS).  The last CTA turns the
// partials into this rank's sums, writes d(gamma), d(beta), and under SyncBN pushes the
// sums to every rank (the all_reduce of _functions.py:159).
struct BnBwdArgs {
  const void* y;      // conv output [B,C,H,H]
  const void* dpool;  // gradient at the pool output [B,C,H/2,H/2]
  const float* params;
  BnStat* st;
  float* bpart;       // [C][NS][2]
  float* gdirect;     // [N_PAD] fp32 gradients written without partials (BN gamma/beta)
  DevState* ds;
  CommDev comm;
  int B, per_split, layer, sync;
  int defer, nsplit;  // defer = 1: pass 1 only writes bpart; the consumer of the sums reduces the nsplit rows
                      // defer = 2: same, bpart holds tagged words and the consumer's first CTA publishes the sums
                      // before it waits for the grid dependency (as bn_forward_publish does)
  // pass 2 only
  void* dy;           // [B,C,H,H] gradient at the conv output
  void* dy_planes;    // optional second copy in the padded-plane layout (conv2_tc.cuh), layer 2 only
  float* gpart;       // conv-bias gradient partial rows
};

template <typename T, int C, int H>
__global__ void __launch_bounds__(256) k_bn_bwd_reduce(BnBwdArgs a) {
  if (threadIdx.x == 0) TRACE((C == 32 ? 7 : 11) * 16 + 0);
  constexpr int HP = H / 2;
  constexpr int OFF_G = (C == 16) ? OFF_G1 : OFF_G2, OFF_BE = (C == 16) ? OFF_BE1 : OFF_BE2;
  __shared__ float red[8][2];
  const int tid = threadIdx.x, c = blockIdx.x;
  const int n0 = blockIdx.y * a.per_split, n1 = min(a.B, n0 + a.per_split);
  const T* __restrict__ y = reinterpret_cast<const T*>(a.y);
  const T* __restrict__ dp = reinterpret_cast<const T*>(a.dpool);
  const float mu = a.st->mean[c], is = a.st->invstd[c], g = a.params[OFF_G + c], b = a.params[OFF_BE + c];
  const unsigned int ll_tag = (unsigned int)a.ds->epoch + 1u;
  pdl_sync();  // predecessor produced dpool; the statistics above come from an earlier kernel
  if (threadIdx.x == 0) TRACE((C == 32 ? 7 : 11) * 16 + 1);
  float sdy = 0.f, sdyx = 0.f;
  const int total = max(0, n1 - n0) * HP * HP;
  // four windows per iteration, every load issued (unconditionally) before any is used: enough bytes in
  // flight per thread to run at HBM speed at large batch
  int e0 = tid;
  for (; e0 + 768 < total; e0 += 1024) {
    float v[4][4], d[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int e = e0 + u * 256;
      const int n = n0 + e / (HP * HP), w = e % (HP * HP), ph = w / HP, pw = w % HP;
      const T* src = y + (((size_t)n * C + c) * H + 2 * ph) * H + 2 * pw;
      load2<T>(src, v[u][0], v[u][1]);
      load2<T>(src + H, v[u][2], v[u][3]);
      d[u] = ldf<T>(&dp[((size_t)n * C + c) * HP * HP + w]);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      bool live;
      float win;
      pool_winner<T>(v[u], mu, is, g, b, live, win);
      if (live) {
        sdy += d[u];
        sdyx = fmaf(d[u], win - mu, sdyx);
      }
    }
  }
  for (; e0 < total; e0 += 256) {  // remainder (and the whole loop at small batch): one window at a time
    const int n = n0 + e0 / (HP * HP), w = e0 % (HP * HP), ph = w / HP, pw = w % HP;
    const T* src = y + (((size_t)n * C + c) * H + 2 * ph) * H + 2 * pw;
    float v[4];
    load2<T>(src, v[0], v[1]);
    load2<T>(src + H, v[2], v[3]);
    const float d = ldf<T>(&dp[((size_t)n * C + c) * HP * HP + w]);
    bool live;
    float win;
    pool_winner<T>(v, mu, is, g, b, live, win);
    if (live) {
      sdy += d;
      sdyx = fmaf(d, win - mu, sdyx);
    }
  }
  sdy = warp_sum(sdy);
  sdyx = warp_sum(sdyx);
  if ((tid & 31) == 0) { red[tid >> 5][0] = sdy; red[tid >> 5][1] = sdyx; }
  __syncthreads();
  if (tid < 2) {
    float v = 0.f;
    for (int w = 0; w < 8; ++w) v += red[w][tid];
    store_partial(a.bpart, ((size_t)c * gridDim.y + blockIdx.y) * 2 + tid, v, a.defer == 2, ll_tag);
  }
  if (threadIdx.x == 0) TRACE((C == 32 ? 7 : 11) * 16 + 15);
  if (a.defer) return;
  if (last_cta_done(&a.ds->ticket[2 + a.layer], gridDim.x * gridDim.y)) {
    const bool xchg = a.sync && a.comm.world > 1;
    if (tid < C) {
      double s0 = 0.0, s1 = 0.0;
#pragma unroll 8
      for (int s = 0; s < (int)gridDim.y; ++s) {
        s0 += (double)__ldcg(&a.bpart[((size_t)tid * gridDim.y + s) * 2]);
        s1 += (double)__ldcg(&a.bpart[((size_t)tid * gridDim.y + s) * 2 + 1]);
      }
      const float f0 = (float)s0, f1 = (float)s1;
      a.st->lsum_dy[tid] = f0;
      a.st->lsum_dy_xmu[tid] = f1;
      a.gdirect[OFF_G + tid] = f1 * a.st->invstd[tid];  // d(gamma), local: DDP averages it
      a.gdirect[OFF_BE + tid] = f0;                      // d(beta)
      if (xchg) {
        const unsigned int tag = (unsigned int)a.ds->epoch + 1u;
        for (int q = 0; q < a.comm.world; ++q) {
          llword* slot = a.comm.peer[q]->ll_bn_bwd[a.layer][a.comm.rank];
          ll_store(&slot[tid], f0, tag);
          ll_store(&slot[C + tid], f1, tag);
        }
      }
    }
  }
}

// Global (all-rank) backward sums for the consumer of pass 1: k1 = sum_dy / N,
// k2 = invstd^2 * sum_dy_xmu / N  (`batch_norm_backward_elemt`, _functions.py:170).
// `after_wait()` runs right after the grid dependency resolved and before the sums are polled: the place for the
// caller's loads of what the direct predecessor wrote, so that they travel together with the polls.
// `consumer_only`: the calling grid is not the designated consumer of pass 1 (a side-branch kernel that also needs the
// sums): none of its CTAs reduces or publishes anything, all of them poll the published words.
template <int C, typename AfterWait>
__device__ void bn_backward_sums(const BnBwdArgs& a, float* s_k1, float* s_k2, AfterWait&& after_wait, bool consumer_only = false) {
  constexpr int OFF_G = (C == 16) ? OFF_G1 : OFF_G2, OFF_BE = (C == 16) ? OFF_BE1 : OFF_BE2;
  __shared__ float s_loc[2 * C];
  const int c = threadIdx.x;
  const bool xchg = a.sync && a.comm.world > 1;
  const bool first = !consumer_only && blockIdx.x == 0 && blockIdx.y == 0;
  // The forward statistics may have been written by the DIRECT predecessor (BN2 inside the Linear+CE kernel): they
  // are read only after the grid dependency resolved, together with the polls.
  float n_tot = 0.f, inv_c = 0.f;
  const unsigned int tag = (unsigned int)a.ds->epoch + 1u;
  // this rank's sums from the pass-1 partials: thread = (quarter of the rows, channel, which sum), so the
  // loads overlap; quarters are combined in fixed order.  CTA (0,0) also publishes d(gamma), d(beta) and
  // pushes the sums to the ranks that need them
  __shared__ float s_q[4][2 * C];
  auto local_sums = [&](bool ll) {
    if (c < 8 * C) {
      const int qd = c / (2 * C), ch = (c % (2 * C)) >> 1, which = c & 1;
      double acc = 0.0;
      if (ll) {  // tagged rows (at most 64, i.e. 16 per thread): polled until they carry this step's tag
        const llword* rows = reinterpret_cast<const llword*>(a.bpart);
        float v[16];
        long long t0 = 0;
        for (int spin = 0;; ++spin) {
          llword w[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const int sp = qd + 4 * i;
            w[i] = sp < a.nsplit ? ll_ld(&rows[((size_t)ch * a.nsplit + sp) * 2 + which]) : ((llword)tag << 32);
          }
          bool ok = true;
#pragma unroll
          for (int i = 0; i < 16; ++i) ok = ok && ((unsigned int)(w[i] >> 32) == tag);
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] = __uint_as_float((unsigned int)w[i]);
          if (ok) break;
          if (spin == 64) t0 = clock64();
          if (spin > 64 && spin_expired(t0, a.ds)) break;
        }
#pragma unroll
        for (int i = 0; i < 16; ++i) acc += (double)v[i];  // rows past nsplit contribute +0.0
      } else {
#pragma unroll 8
        for (int sp = qd; sp < a.nsplit; sp += 4) acc += (double)__ldcg(&a.bpart[((size_t)ch * a.nsplit + sp) * 2 + which]);
      }
      s_q[qd][which * C + ch] = (float)acc;
    }
    __syncthreads();
    if (c < 2 * C) s_loc[c] = ((s_q[0][c] + s_q[1][c]) + s_q[2][c]) + s_q[3][c];
    __syncthreads();
  };
  auto publish = [&](bool self_too) {
    if (c < C) {
      for (int q = 0; q < a.comm.world; ++q) {  // the all_reduce of _functions.py:159 as pushed tagged words
        if (q == a.comm.rank ? !self_too : !xchg) continue;
        llword* slot = a.comm.peer[q]->ll_bn_bwd[a.layer][a.comm.rank];
        ll_store(&slot[c], s_loc[c], tag);
        ll_store(&slot[C + c], s_loc[C + c], tag);
      }
    }
  };
  const bool polled = a.defer == 2;
  if (polled && first) {  // before the wait: the rows are picked up while the pass-1 grid is still draining
    local_sums(true);
    publish(true);
  }
  pdl_sync();  // predecessor: pass 1 (bpart)
  after_wait();
  if (c < C) { n_tot = a.st->n_total; inv_c = a.st->invstd[c]; }
  if (a.defer == 1) {
    local_sums(false);
    if (first) publish(xchg);
  }
  if (first && a.defer != 0 && c < C) {  // this rank's d(gamma), d(beta): DDP averages them (s_loc: CTA (0,0)'s sums)
    a.gdirect[OFF_G + c] = s_loc[C + c] * inv_c;
    a.gdirect[OFF_BE + c] = s_loc[c];
  }
  if (c < C) {
    float s0, s1;
    if (polled && !xchg && first) {  // single rank, the publishing CTA: it has the sums already
      s0 = s_loc[c];
      s1 = s_loc[C + c];
    } else if (polled && !xchg) {  // single rank: the words CTA (0,0) published
      float got[2];
      const int idx[2] = {c, C + c};
      ll_gather<2, 1>(&a.comm.self->ll_bn_bwd[a.layer][a.comm.rank][0], 64, idx, 0, 1, tag, got, a.ds);
      s0 = got[0]; s1 = got[1];
    } else if (xchg) {
      s0 = 0.f; s1 = 0.f;
      float got[MAX_WORLD * 2];
      const int idx[2] = {c, C + c};
      ll_gather<2>(&a.comm.self->ll_bn_bwd[a.layer][0][0], 64, idx, 0, a.comm.world, tag, got, a.ds);
#pragma unroll
      for (int r = 0; r < MAX_WORLD; ++r) {  // fixed rank order
        if (r < a.comm.world) { s0 += got[r * 2]; s1 += got[r * 2 + 1]; }
      }
    } else if (a.defer) {
      s0 = s_loc[c];
      s1 = s_loc[C + c];
    } else {
      s0 = a.st->lsum_dy[c];
      s1 = a.st->lsum_dy_xmu[c];
    }
    s_k1[c] = s0 / n_tot;
    s_k2[c] = inv_c * inv_c * s1 / n_tot;
  }
  __syncthreads();
}

template <int C>
__device__ __forceinline__ void bn_backward_sums(const BnBwdArgs& a, float* s_k1, float* s_k2) {
  bn_backward_sums<C>(a, s_k1, s_k2, [] {});
}

// BN backward, pass 2 for layer 2: d(conv2 out) materialised (it feeds both conv2 dgrad and
// wgrad); also the conv2 bias gradient partials (sum of the rounded values).  grid (32, NS).
template <typename T>
__global__ void __launch_bounds__(256) k_bn2_bwd_apply(BnBwdArgs a) {
  if (threadIdx.x == 0) TRACE(9 * 16 + 0);
  constexpr int C = 32, H = 14, HP = 7;
  __shared__ float s_k1[C], s_k2[C];
  __shared__ float red[8];
  const int tid = threadIdx.x, c = blockIdx.x;
  const int n0 = blockIdx.y * a.per_split, n1 = min(a.B, n0 + a.per_split);
  const T* __restrict__ y = reinterpret_cast<const T*>(a.y);
  const T* __restrict__ dp = reinterpret_cast<const T*>(a.dpool);
  T* __restrict__ dy = reinterpret_cast<T*>(a.dy);
  const int total = max(0, n1 - n0) * HP * HP;
  // The conv output of this thread's first window is older than the predecessor's predecessor (a kernel is only
  // launched once every CTA of the one before it has passed ITS wait), so it is fetched before the wait inside
  // bn_backward_sums; the incoming gradient and BN2's statistics may come from the direct predecessor (pass 1 and
  // BN2 itself fused into the Linear+CE kernel) and are fetched right after the wait, together with the polls
  const float g = a.params[OFF_G2 + c], b = a.params[OFF_BE2 + c];
  float mu = 0.f, is = 1.f;  // BN2's statistics may come from the direct predecessor too: fetched after the wait
  float pv[4] = {0.f, 0.f, 0.f, 0.f}, pd = 0.f;
  if (tid < total) {
    const int n = n0 + tid / (HP * HP), w = tid % (HP * HP), ph = w / HP, pw = w % HP;
    const size_t off = (((size_t)n * C + c) * H + 2 * ph) * H + 2 * pw;
    load2<T>(y + off, pv[0], pv[1]);
    load2<T>(y + off + H, pv[2], pv[3]);
  }
  bn_backward_sums<C>(a, s_k1, s_k2, [&] {
    mu = a.st->mean[c];
    is = a.st->invstd[c];
    if (tid < total) pd = ldf<T>(&dp[((size_t)(n0 + tid / (HP * HP)) * C + c) * HP * HP + tid % (HP * HP)]);
  });
  if (threadIdx.x == 0) TRACE(9 * 16 + 1);
  const float k1 = s_k1[c], k2 = s_k2[c], k3 = is * g;
  float bsum = 0.f;
  for (int e = tid; e < total; e += 256) {
    const int n = n0 + e / (HP * HP), w = e % (HP * HP), ph = w / HP, pw = w % HP;
    const size_t off = (((size_t)n * C + c) * H + 2 * ph) * H + 2 * pw;
    float v[4], dd;
    if (e == tid) {
      v[0] = pv[0]; v[1] = pv[1]; v[2] = pv[2]; v[3] = pv[3]; dd = pd;
    } else {
      load2<T>(y + off, v[0], v[1]);
      load2<T>(y + off + H, v[2], v[3]);
      dd = ldf<T>(&dp[((size_t)n * C + c) * HP * HP + w]);
    }
    bool live;
    const int arg = pool_winner<T>(v, mu, is, g, b, live);
    const float d = live ? dd : 0.f;
    float o[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float da = (live && i == arg) ? d : 0.f;
      o[i] = rnd<T>((da - k1 - (v[i] - mu) * k2) * k3);
      bsum += o[i];
    }
    store2<T>(dy + off, o[0], o[1]);
    store2<T>(dy + off + H, o[2], o[3]);
    if (a.dy_planes != nullptr) {
      T* pl = reinterpret_cast<T*>(a.dy_planes) + (((size_t)(n * 4 + (c >> 3)) * 320 + (2 * ph + 3) * 16 + 2 * pw) * 8 + (c & 7));
      stf<T>(pl, o[0]); stf<T>(pl + 8, o[1]); stf<T>(pl + 128, o[2]); stf<T>(pl + 136, o[3]);
    }
  }
  if (threadIdx.x == 0) TRACE(9 * 16 + 2);
  bsum = warp_sum(bsum);
  if ((tid & 31) == 0) red[tid >> 5] = bsum;
  __syncthreads();
  if (tid == 0) {
    float v = 0.f;
    for (int w = 0; w < 8; ++w) v += red[w];
    a.gpart[(size_t)blockIdx.y * N_PAD + OFF_B2 + c] = v;
  }
  if (threadIdx.x == 0) TRACE(9 * 16 + 15);
}

// Same, for the tensor-core path: d(conv2 out) is written ONLY in the padded-plane layout
// (conv2_tc.cuh), thread = one pooling window x 8 channels -> four 16-byte stores.  grid (4 channel
// groups, NS batch splits).
template <typename T>
__global__ void __launch_bounds__(256) k_bn2_bwd_apply_planes(BnBwdArgs a) {
  constexpr int C = 32, H = 14, HP = 7;
  __shared__ float s_k1[C], s_k2[C];
  __shared__ float s_c[8][4];  // this group's channels: mean, invstd, gamma, beta
  __shared__ float red[8][8];
  const int tid = threadIdx.x, g = blockIdx.x;
  float c0 = 0.f, c1 = 0.f, c2 = 0.f, c3 = 0.f;
  if (tid < 8) {  // independent of the predecessor: issued before the wait inside bn_backward_sums
    const int c = g * 8 + tid;
    c0 = a.st->mean[c]; c1 = a.st->invstd[c]; c2 = a.params[OFF_G2 + c]; c3 = a.params[OFF_BE2 + c];
  }
  bn_backward_sums<C>(a, s_k1, s_k2);
  if (tid < 8) { s_c[tid][0] = c0; s_c[tid][1] = c1; s_c[tid][2] = c2; s_c[tid][3] = c3; }
  __syncthreads();
  const int n0 = blockIdx.y * a.per_split, n1 = min(a.B, n0 + a.per_split);
  const T* __restrict__ y = reinterpret_cast<const T*>(a.y);
  const T* __restrict__ dp = reinterpret_cast<const T*>(a.dpool);
  T* __restrict__ pl = reinterpret_cast<T*>(a.dy_planes);
  float bsum[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) bsum[j] = 0.f;
  const int total = max(0, n1 - n0) * HP * HP;
  for (int e = tid; e < total; e += 256) {
    const int n = n0 + e / (HP * HP), w = e % (HP * HP), ph = w / HP, pw = w % HP;
    float o[4][8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int c = g * 8 + j;
      const size_t off = (((size_t)n * C + c) * H + 2 * ph) * H + 2 * pw;
      float v[4];
      load2<T>(y + off, v[0], v[1]);
      load2<T>(y + off + H, v[2], v[3]);
      const float mu = s_c[j][0], is = s_c[j][1], gm = s_c[j][2];
      bool live;
      const int arg = pool_winner<T>(v, mu, is, gm, s_c[j][3], live);
      const float d = live ? ldf<T>(&dp[((size_t)n * C + c) * HP * HP + w]) : 0.f;
      const float k1 = s_k1[c], k2 = s_k2[c], k3 = is * gm;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float da = (live && i == arg) ? d : 0.f;
        o[i][j] = rnd<T>((da - k1 - (v[i] - mu) * k2) * k3);
        bsum[j] += o[i][j];
      }
    }
    T* dst = pl + (((size_t)(n * 4 + g) * 320 + (2 * ph + 3) * 16 + 2 * pw) * 8);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      T* d2 = dst + ((i >> 1) * 16 + (i & 1)) * 8;
      store4<T>(d2, o[i][0], o[i][1], o[i][2], o[i][3]);
      store4<T>(d2 + 4, o[i][4], o[i][5], o[i][6], o[i][7]);
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float v = warp_sum(bsum[j]);
    if ((tid & 31) == 0) red[tid >> 5][j] = v;
  }
  __syncthreads();
  if (tid < 8) {
    float v = 0.f;
    for (int w = 0; w < 8; ++w) v += red[w][tid];
    a.gpart[(size_t)blockIdx.y * N_PAD + OFF_B2 + g * 8 + tid] = v;
  }
}

// ---------------------------------------------------------------------------------------
// conv2 weight gradient: dW[co][ci][kh][kw] = sum_{n,h,w} dy2[n,co,h,w] * p1[n,ci,h+kh-2,w+kw-2]
// grid (GX images-strided, 4 groups of 8 output channels), 160 threads = 16 ci x 5 kh x 2
// quads of co; each thread owns a 5 (kw) x 4 (co) register tile over all its images.
struct Conv2WgradArgs {
  const void* p1;   // [B,16,14,14]
  const void* dy2;  // [B,32,14,14]
  float* gpart;     // [rows = gridDim.x][N_PAD]
  int B;
};

template <typename T>
__global__ void __launch_bounds__(160) k_conv2_wgrad(Conv2WgradArgs a) {
  __shared__ __align__(16) float in_s[16 * 18 * C14_RS];
  __shared__ __align__(16) float dy_s[196 * 8];
  const int tid = threadIdx.x, cog = blockIdx.y;
  const int coq = tid & 1, r = tid >> 1, kh = r % 5, ci = r / 5;
  const T* __restrict__ p1 = reinterpret_cast<const T*>(a.p1);
  const T* __restrict__ dy2 = reinterpret_cast<const T*>(a.dy2);
  for (int i = tid; i < 16 * 18 * C14_RS; i += 160) in_s[i] = 0.f;
  float acc[5][4];
#pragma unroll
  for (int k = 0; k < 5; ++k)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[k][j] = 0.f;

  for (int n = blockIdx.x; n < a.B; n += gridDim.x) {
    __syncthreads();
    for (int i = tid; i < 16 * 196; i += 160) {
      const int c = i / 196, rr = (i % 196) / 14, cc = i % 14;
      in_s[(c * 18 + rr + 2) * C14_RS + cc + 2] = ldf<T>(&p1[(size_t)n * 16 * 196 + i]);
    }
    for (int i = tid; i < 8 * 196; i += 160) {
      const int cl = i / 196, pix = i % 196;
      dy_s[pix * 8 + cl] = ldf<T>(&dy2[((size_t)n * 32 + cog * 8 + cl) * 196 + pix]);
    }
    __syncthreads();
    for (int h = 0; h < 14; ++h) {
      float xr[18];
      const float* xrow = in_s + (ci * 18 + h + kh) * C14_RS;
#pragma unroll
      for (int c = 0; c < 18; ++c) xr[c] = xrow[c];
#pragma unroll
      for (int w = 0; w < 14; ++w) {
        const float4 d = *reinterpret_cast<const float4*>(&dy_s[(h * 14 + w) * 8 + coq * 4]);
#pragma unroll
        for (int kw = 0; kw < 5; ++kw) {
          acc[kw][0] = fmaf(xr[w + kw], d.x, acc[kw][0]);
          acc[kw][1] = fmaf(xr[w + kw], d.y, acc[kw][1]);
          acc[kw][2] = fmaf(xr[w + kw], d.z, acc[kw][2]);
          acc[kw][3] = fmaf(xr[w + kw], d.w, acc[kw][3]);
        }
      }
    }
  }
  float* row = a.gpart + (size_t)blockIdx.x * N_PAD;
#pragma unroll
  for (int j = 0; j < 4; ++j)
#pragma unroll
    for (int kw = 0; kw < 5; ++kw)
      row[OFF_W2 + ((cog * 8 + coq * 4 + j) * 16 + ci) * 25 + kh * 5 + kw] = acc[kw][j];
}

// ---------------------------------------------------------------------------------------
// BN1 backward pass 2 fused with the conv1 weight / bias gradient (conv1's input needs no
// gradient, so d(conv1 out) lives only in shared memory):
//   dW1[co][kh][kw] = sum_{n,h,w} dy1[n,co,h,w] * x[n,h+kh-2,w+kw-2],  db1[co] = sum dy1
// CTA = one image x 4 channels (grid.y = 4), 160 threads.  Phase A: five pooling windows per thread
// -> dy1 in shared memory (their loads are issued before the wait for the BN sums).  Phase B: thread = (channel, kh, 4 of the 28 rows)
// with a 5-wide kw register tile.
struct Conv1WgradArgs {
  BnBwdArgs bn;    // y = y1, dpool = d(pool1 out), layer = 0
  BatchSrc x;      // the input batch (same source as conv1 forward)
};

constexpr int C1W_THREADS = 160;

template <typename T>
__global__ void __launch_bounds__(C1W_THREADS) k_conv1_wgrad(Conv1WgradArgs a) {
  if (threadIdx.x == 0) TRACE(12 * 16 + 0);
  __shared__ __align__(16) float dy_s[4 * 784];
  __shared__ __align__(16) float xs[32 * C1_XS];
  __shared__ float red[140][6];
  __shared__ float s_k1[16], s_k2[16];
  __shared__ float s_cst[4][6];  // per channel of this CTA: mean, invstd, gamma, beta, k1, k2
  const int tid = threadIdx.x, cg = blockIdx.y;
  float c_mu = 0.f, c_is = 0.f, c_g = 0.f, c_b = 0.f;
  if (tid < 4) {  // fetched before the sums so the latencies overlap
    const int c = cg * 4 + tid;
    c_mu = a.bn.st->mean[c]; c_is = a.bn.st->invstd[c];
    c_g = a.bn.params[OFF_G1 + c]; c_b = a.bn.params[OFF_BE1 + c];
  }
  for (int i = tid; i < 32 * C1_XS; i += C1W_THREADS) xs[i] = 0.f;
  if (tid < 4) { s_cst[tid][0] = c_mu; s_cst[tid][1] = c_is; s_cst[tid][2] = c_g; s_cst[tid][3] = c_b; }
  const T* __restrict__ y1 = reinterpret_cast<const T*>(a.bn.y);
  const T* __restrict__ dp1 = reinterpret_cast<const T*>(a.bn.dpool);
  const int rs = tid % 7, kh = (tid / 7) % 5, col = tid / 35;  // valid for tid < 140
  float acc[5] = {0.f, 0.f, 0.f, 0.f, 0.f}, bsum = 0.f;
  // Phase A, part 1 for one image: the input tile into shared memory and this thread's five pooling windows
  // (conv1 output + incoming gradient) into registers -- all loads issued together
  float wv[5][4], wd[5];
  auto fetch_grad = [&](int n) {  // d(pool1 out): may come from the direct predecessor, so only after the wait
#pragma unroll
    for (int it = 0; it < 5; ++it) {
      const int e = min(tid + it * C1W_THREADS, 783);
      wd[it] = ldf<T>(&dp1[((size_t)n * 16 + cg * 4 + e / 196) * 196 + e % 196]);
    }
  };
  auto fetch = [&](int n) {
    for (int i = tid; i < 196; i += C1W_THREADS) {
      const float4 v = load_pixels4(a.x, n, i);
      float* d = &xs[((4 * i) / 28 + 2) * C1_XS + (4 * i) % 28 + 2];
      d[0] = rnd<T>(v.x); d[1] = rnd<T>(v.y); d[2] = rnd<T>(v.z); d[3] = rnd<T>(v.w);
    }
#pragma unroll
    for (int it = 0; it < 5; ++it) {
      const int e = min(tid + it * C1W_THREADS, 783);
      const int cl = e / 196, w = e % 196, ph = w / 14, pw = w % 14, c = cg * 4 + cl;
      const size_t off = (((size_t)n * 16 + c) * 28 + 2 * ph) * 28 + 2 * pw;
      load2<T>(y1 + off, wv[it][0], wv[it][1]);
      load2<T>(y1 + off + 28, wv[it][2], wv[it][3]);
    }
  };
  // The input and conv1's output do not depend on the predecessor, only on kernels before it, which are complete
  // once this grid runs: the first image is fetched BEFORE waiting for the sums, its incoming gradient right after
  __syncthreads();
  if ((int)blockIdx.x < a.bn.B) fetch(blockIdx.x);
  bn_backward_sums<16>(a.bn, s_k1, s_k2, [&] { if ((int)blockIdx.x < a.bn.B) fetch_grad(blockIdx.x); });
  if (threadIdx.x == 0) TRACE(12 * 16 + 1);
  if (tid < 4) { s_cst[tid][4] = s_k1[cg * 4 + tid]; s_cst[tid][5] = s_k2[cg * 4 + tid]; }

  for (int n = blockIdx.x; n < a.bn.B; n += gridDim.x) {
    __syncthreads();
    if (n != (int)blockIdx.x) { fetch(n); fetch_grad(n); }
#pragma unroll
    for (int it = 0; it < 5; ++it) {
      const int e = tid + it * C1W_THREADS;
      if (e < 784) {
        const int cl = e / 196, w = e % 196, ph = w / 14, pw = w % 14;
        const float mu = s_cst[cl][0], is = s_cst[cl][1], g = s_cst[cl][2];
        bool live;
        const int arg = pool_winner<T>(wv[it], mu, is, g, s_cst[cl][3], live);
        const float d = live ? wd[it] : 0.f;
        const float k1 = s_cst[cl][4], k2 = s_cst[cl][5], k3 = is * g;
        float* dst = dy_s + (cl * 28 + 2 * ph) * 28 + 2 * pw;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float da = (live && i == arg) ? d : 0.f;
          dst[(i >> 1) * 28 + (i & 1)] = rnd<T>((da - k1 - (wv[it][i] - mu) * k2) * k3);
        }
      }
    }
    __syncthreads();
    if (threadIdx.x == 0) TRACE(12 * 16 + 2);
    if (tid < 140) {
#pragma unroll
      for (int hh = 0; hh < 4; ++hh) {
        const int h = hh * 7 + rs;  // lanes of a warp read consecutive rows: bank-conflict free 16-byte loads
        const float* drow = dy_s + (col * 28 + h) * 28;
        const float* xrow = xs + (h + kh) * C1_XS;
#pragma unroll
        for (int q = 0; q < 7; ++q) {
          const float4 d = *reinterpret_cast<const float4*>(drow + 4 * q);
          const float4 u = *reinterpret_cast<const float4*>(xrow + 4 * q);
          const float4 v = *reinterpret_cast<const float4*>(xrow + 4 * q + 4);
          const float xv[8] = {u.x, u.y, u.z, u.w, v.x, v.y, v.z, v.w};
#pragma unroll
          for (int kw = 0; kw < 5; ++kw) {
            acc[kw] = fmaf(d.x, xv[kw], acc[kw]);
            acc[kw] = fmaf(d.y, xv[kw + 1], acc[kw]);
            acc[kw] = fmaf(d.z, xv[kw + 2], acc[kw]);
            acc[kw] = fmaf(d.w, xv[kw + 3], acc[kw]);
          }
          if (kh == 0) bsum += (d.x + d.y) + (d.z + d.w);
        }
      }
    }
  }
  if (threadIdx.x == 0) TRACE(12 * 16 + 3);
  // fixed-order reduction over the 7 row groups
  if (tid < 140) {
#pragma unroll
    for (int kw = 0; kw < 5; ++kw) red[tid][kw] = acc[kw];
    red[tid][5] = bsum;
  }
  __syncthreads();
  float* row = a.bn.gpart + (size_t)blockIdx.x * N_PAD;
  if (tid < 100) {  // (col, kh, kw)
    const int cl = tid / 25, k = tid % 25, kh2 = k / 5, kw2 = k % 5;
    float v = 0.f;
#pragma unroll
    for (int r = 0; r < 7; ++r) v += red[(cl * 5 + kh2) * 7 + r][kw2];
    row[OFF_W1 + (cg * 4 + cl) * 25 + k] = v;
  } else if (tid < 104) {
    const int cl = tid - 100;
    float v = 0.f;
#pragma unroll
    for (int r = 0; r < 7; ++r) v += red[(cl * 5 + 0) * 7 + r][5];
    row[OFF_B1 + cg * 4 + cl] = v;
  }
  if (threadIdx.x == 0) TRACE(12 * 16 + 15);
}

}  // namespace dstep
