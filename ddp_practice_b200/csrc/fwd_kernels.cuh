// Forward kernels of the ConvNet step (direct, CUDA-core path; any batch size).
//   conv1 5x5 p2 (1->16)  + per-channel BN statistics in the epilogue     ddp_main.py:17-18
//   BN apply + ReLU + MaxPool2                                            ddp_main.py:18-20,24-26
//   conv2 5x5 p2 (16->32) + BN statistics (also: its data-gradient form)  ddp_main.py:23
//   Linear(1568,10) + CrossEntropyLoss + its backward, fused               ddp_main.py:28,89,91
#pragma once
#include "common.cuh"

namespace dstep {

template <typename T> __device__ __forceinline__ void store4(T* p, float a, float b, float c, float d);
template <> __device__ __forceinline__ void store4<float>(float* p, float a, float b, float c, float d) {
  *reinterpret_cast<float4*>(p) = make_float4(a, b, c, d);
}
template <> __device__ __forceinline__ void store4<__half>(__half* p, float a, float b, float c, float d) {
  __half2 lo = __floats2half2_rn(a, b), hi = __floats2half2_rn(c, d);
  uint2 u;
  u.x = *reinterpret_cast<unsigned int*>(&lo);
  u.y = *reinterpret_cast<unsigned int*>(&hi);
  *reinterpret_cast<uint2*>(p) = u;
}
template <> __device__ __forceinline__ void store4<__nv_bfloat16>(__nv_bfloat16* p, float a, float b, float c, float d) {
  __nv_bfloat162 lo = __floats2bfloat162_rn(a, b), hi = __floats2bfloat162_rn(c, d);
  uint2 u;
  u.x = *reinterpret_cast<unsigned int*>(&lo);
  u.y = *reinterpret_cast<unsigned int*>(&hi);
  *reinterpret_cast<uint2*>(p) = u;
}

template <typename T> __device__ __forceinline__ void load2(const T* p, float& a, float& b);
template <> __device__ __forceinline__ void load2<float>(const float* p, float& a, float& b) {
  float2 v = *reinterpret_cast<const float2*>(p);
  a = v.x; b = v.y;
}
template <> __device__ __forceinline__ void load2<__half>(const __half* p, float& a, float& b) {
  float2 v = __half22float2(*reinterpret_cast<const __half2*>(p));
  a = v.x; b = v.y;
}
template <> __device__ __forceinline__ void load2<__nv_bfloat16>(const __nv_bfloat16* p, float& a, float& b) {
  float2 v = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(p));
  a = v.x; b = v.y;
}
template <typename T> __device__ __forceinline__ void store2(T* p, float a, float b);
template <> __device__ __forceinline__ void store2<float>(float* p, float a, float b) {
  *reinterpret_cast<float2*>(p) = make_float2(a, b);
}
template <> __device__ __forceinline__ void store2<__half>(__half* p, float a, float b) {
  *reinterpret_cast<__half2*>(p) = __floats2half2_rn(a, b);
}
template <> __device__ __forceinline__ void store2<__nv_bfloat16>(__nv_bfloat16* p, float a, float b) {
  *reinterpret_cast<__nv_bfloat162*>(p) = __floats2bfloat162_rn(a, b);
}

// conv2's filter as ready-made tcgen05 B-operand tiles (conv2_tc.cuh): element i of [2][25][1 KB], gathered from the
// fp32 master weights.  i < 12800: forward [tap][ci/8][co][ci%8]; else data gradient [tap'][co/8][ci][co%8], tap' = 24 - tap.
template <typename T>
__device__ __forceinline__ void w2_tile_gather(const float* __restrict__ params, void* tiles, int i) {
  if (i >= 25600) return;
  const int k8 = i & 7;
  float w;
  if (i < 12800) {
    const int co = (i >> 3) & 31, kc = (i >> 8) & 1, tap = i >> 9;
    w = params[OFF_W2 + (co * 16 + kc * 8 + k8) * 25 + tap];
  } else {
    const int j = i - 12800;
    const int ci = (j >> 3) & 15, kc = (j >> 7) & 3, tp = j >> 9;
    w = params[OFF_W2 + ((kc * 8 + k8) * 16 + ci) * 25 + (24 - tp)];
  }
  reinterpret_cast<T*>(tiles)[i] = Cvt<T>::from_f(w);
}

// ---------------------------------------------------------------------------------------
// Local BN statistics from per-CTA (sum, sum of squares) partials; run by the last CTA of a
// conv kernel.  `batch_norm_stats` (torch/nn/modules/_functions.py:39).  With SyncBN the
// (mean, invstd, count) triple is pushed to every rank (the all_gather of _functions.py:74).
template <int C>
__device__ void finalize_fwd_stats(const float* part, int rows, float count, float eps, BnStat* st, int layer,
                                   const CommDev& comm, DevState* ds, int sync) {
  // all threads of the CTA take part: column = (channel, statistic), several row lanes per column,
  // so the loads from the partial buffer overlap instead of forming one long dependent chain
  __shared__ double fin[512];
  constexpr int COLS = 2 * C;
  const int t = threadIdx.x;
  const int lanes = min((int)blockDim.x / COLS, 512 / COLS);
  const int col = t % COLS, lane = t / COLS;
  if (lane < lanes) {
    double acc = 0.0;
#pragma unroll 4
    for (int r = lane; r < rows; r += lanes) acc += (double)__ldcg(&part[(size_t)r * COLS + col]);
    fin[lane * COLS + col] = acc;
  }
  __syncthreads();
  const int c = t;
  const bool xchg = sync && comm.world > 1;
  if (c < C) {
    double s = 0.0, ss = 0.0;
    for (int l = 0; l < lanes; ++l) {
      s += fin[l * COLS + 2 * c];
      ss += fin[l * COLS + 2 * c + 1];
    }
    double mean = s / (double)count;
    double var = ss / (double)count - mean * mean;
    if (var < 0.0) var = 0.0;
    float invstd = (float)(1.0 / sqrt(var + (double)eps));
    st->lmean[c] = (float)mean;
    st->linvstd[c] = invstd;
    st->lvar_unbiased[c] = (float)(var * ((double)count / ((double)count - 1.0)));
    if (xchg) {
      const unsigned int tag = (unsigned int)ds->epoch + 1u;
      for (int q = 0; q < comm.world; ++q) {
        llword* slot = comm.peer[q]->ll_bn_fwd[layer][comm.rank];
        ll_store(&slot[c], (float)mean, tag);
        ll_store(&slot[C + c], invstd, tag);
        if (c == 0) ll_store(&slot[3 * C], count, tag);
      }
    }
  }
}

// ---------------------------------------------------------------------------------------
// conv1: x fp32 [B,1,28,28] -> y1 T [B,16,28,28].  CTA = one image x 4 output channels
// (grid.y = 4 channel groups, so a batch of 32 already fills 128 SMs), thread = one 2x4 pixel
// tile x 4 channels, grid-stride over images.  K = 25 makes this a bandwidth kernel: the image
// is staged once in shared memory, every output is written once with 8/16-byte stores, the BN
// reduction rides in the epilogue.
struct Conv1Args {
  BatchSrc x;
  const float* params;
  void* y1;
  float* part;  // [gridDim.x][16][2]
  int defer;    // 1: leave the partial sums to the consumer (bn_forward_stats), no last-CTA finalisation;
                // 2: same, and `part` holds tagged 8-byte words (store_partial)
  BnStat* st;
  DevState* ds;
  CommDev comm;
  int B, dbl, sync;
  float eps;
  // small-batch graph: CTAs with blockIdx.x >= conv_ctas rebuild the conv2 filter tiles (conv2_tc.cuh) instead of
  // convolving, so the tiles are ready when this kernel completes and conv2 needs no extra graph branch / join
  void* w2_tiles;
  int conv_ctas;
};

constexpr int C1_XS = 36;       // padded row stride of the staged image
constexpr int C1_THREADS = 128; // 98 active tiles
constexpr int C1_GROUPS = 4;    // gridDim.y

template <typename T>
__global__ void __launch_bounds__(C1_THREADS) k_conv1_fwd(Conv1Args a) {
  if (threadIdx.x == 0) TRACE(1 * 16 + 0);
  __shared__ __align__(16) float xs[32 * C1_XS];
  __shared__ __align__(16) float ws[25 * 4];
  __shared__ float bs[4];
  __shared__ float red[4][8];
  const int tid = threadIdx.x, cg = blockIdx.y;
  T* __restrict__ y1 = reinterpret_cast<T*>(a.y1);
  // First kernel of the step.  Alone in its graph it has nothing to wait for; in a multi-step graph
  // (ddpstep_train_steps) its programmatic predecessor is the PREVIOUS step's tail kernel, which writes the parameters
  // and the step counter -- so both are read only after the wait.
  pdl_sync();
  const unsigned int ll_tag = (unsigned int)a.ds->epoch + 1u;
  if (a.w2_tiles != nullptr && (int)blockIdx.x >= a.conv_ctas) {
    w2_tile_gather<T>(a.params, a.w2_tiles, (((int)blockIdx.x - a.conv_ctas) * C1_GROUPS + cg) * C1_THREADS + tid);
    return;
  }

  for (int i = tid; i < 32 * C1_XS; i += C1_THREADS) xs[i] = 0.f;
  if (tid < 100) {  // ws[k][j] <- rnd(W1[cg*4+j][k])
    const int j = tid / 25, k = tid % 25;
    ws[k * 4 + j] = rnd<T>(a.params[OFF_W1 + (cg * 4 + j) * 25 + k]);
  }
  if (tid < 4) bs[tid] = rnd<T>(a.params[OFF_B1 + cg * 4 + tid]);

  const bool active = tid < 98;
  const int tr = tid / 7, tc = tid % 7;
  float s[4], ss[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) s[j] = ss[j] = 0.f;

  for (int n = blockIdx.x; n < a.B; n += a.conv_ctas) {  // conv_ctas, not gridDim.x: the grid may carry extra filter-tile CTAs
    __syncthreads();
    for (int i = tid; i < 196; i += C1_THREADS) {  // 784 floats as 196 float4
      const float4 v = load_pixels4(a.x, n, i);
      const int r = (4 * i) / 28, c = (4 * i) % 28;
      float* d = &xs[(r + 2) * C1_XS + c + 2];
      d[0] = rnd<T>(v.x); d[1] = rnd<T>(v.y); d[2] = rnd<T>(v.z); d[3] = rnd<T>(v.w);
    }
    __syncthreads();
    if (threadIdx.x == 0) TRACE(1 * 16 + 1);
    if (!active) continue;
    float acc[2][4][4];
#pragma unroll
    for (int r = 0; r < 2; ++r)
#pragma unroll
      for (int c = 0; c < 4; ++c)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[r][c][j] = 0.f;
    // fully unrolled: a rolled kh loop (8x less code) was measured 20% SLOWER, so instruction fetch is not what
    // limits this body; one warp per scheduler issuing 3-register FFMAs is
#pragma unroll
    for (int kh = 0; kh < 5; ++kh) {
      float in[2][8];
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        const float4 u = *reinterpret_cast<const float4*>(&xs[(2 * tr + r + kh) * C1_XS + 4 * tc]);
        const float4 v = *reinterpret_cast<const float4*>(&xs[(2 * tr + r + kh) * C1_XS + 4 * tc + 4]);
        in[r][0] = u.x; in[r][1] = u.y; in[r][2] = u.z; in[r][3] = u.w;
        in[r][4] = v.x; in[r][5] = v.y; in[r][6] = v.z; in[r][7] = v.w;
      }
#pragma unroll
      for (int kw = 0; kw < 5; ++kw) {
        const float4 w0 = *reinterpret_cast<const float4*>(&ws[(kh * 5 + kw) * 4]);
        const float w[4] = {w0.x, w0.y, w0.z, w0.w};
#pragma unroll
        for (int r = 0; r < 2; ++r)
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            const float v = in[r][c + kw];
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[r][c][j] = fmaf(v, w[j], acc[r][c][j]);
          }
      }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int ch = cg * 4 + j;
      const float b = bs[j];
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        float v[4];
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          v[c] = conv_out<T>(acc[r][c][j], b, a.dbl);
          s[j] += v[c];
          ss[j] = fmaf(v[c], v[c], ss[j]);
        }
        store4<T>(&y1[(((size_t)n * 16 + ch) * 28 + 2 * tr + r) * 28 + 4 * tc], v[0], v[1], v[2], v[3]);
      }
    }
  }
  if (a.part == nullptr) return;  // eval: running statistics are used, none are gathered
  if (threadIdx.x == 0) TRACE(1 * 16 + 2);
  __syncthreads();
  const int warp = tid >> 5, lane = tid & 31;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    float v = warp_sum(s[j]), q = warp_sum(ss[j]);
    if (lane == 0) { red[warp][j] = v; red[warp][4 + j] = q; }
  }
  __syncthreads();
  if (tid < 8) {  // tid = stat*4 + j
    const int stat = tid >> 2, j = tid & 3;
    const float v = ((red[0][tid] + red[1][tid]) + red[2][tid]) + red[3][tid];
    store_partial(a.part, ((size_t)blockIdx.x * 16 + cg * 4 + j) * 2 + stat, v, a.defer == 2, ll_tag);
  }
  if (threadIdx.x == 0) TRACE(1 * 16 + 15);
  if (a.defer) return;
  if (last_cta_done(&a.ds->ticket[0], a.conv_ctas * gridDim.y))  // the filter-tile CTAs returned long ago
    finalize_fwd_stats<16>(a.part, a.conv_ctas, (float)a.B * 784.f, a.eps, a.st, 0, a.comm, a.ds, a.sync);
}

// ---------------------------------------------------------------------------------------
// BN apply + ReLU + MaxPool 2x2.  Every CTA first obtains the statistics to normalise with:
// train/W=1: this rank's; train/SyncBN: Chan-merge of all ranks' (mean, invstd, count) in rank
// order (`batch_norm_gather_stats_with_counts`, _functions.py:106); eval: running buffers.
// CTA 0 also performs the running-statistics update and num_batches_tracked += 1.
struct BnPoolArgs {
  const void* y;
  void* p;
  void* p_planes;  // optional second copy in the padded-plane layout the tcgen05 conv2 reads (conv2_tc.cuh)
  const float* params;
  BnStat* st;
  float* bnbuf;
  long long* nbt;
  DevState* ds;
  CommDev comm;
  int B, layer, eval, sync;
  float momentum, eps;
  // deferred finalisation (small grids): the conv kernel only wrote its per-CTA partial sums and
  // THIS kernel's CTAs reduce them in their prologue -- no fence / ticket / serial last-CTA tail in
  // the producer.  part == nullptr: the producer finalised (large grids) and st->l* are valid.
  const float* part;
  int rows;
  int ll;  // `part` holds tagged words: CTA 0 publishes the local statistics early (bn_forward_publish), the others poll
};

// Local statistics from the producer's partial rows (column = (channel, stat), several row lanes per column, sums in
// double, divisions / square root in fp32 as batch_norm_stats does).  Called by every thread of a 256-thread CTA.
// LL = true: the rows are tagged words that are polled until they carry this step's tag.
template <int C, bool LL, int BLOCK = 256>
__device__ void bn_local_stats(const BnPoolArgs& a, float local_count, unsigned int tag, float& lmean, float& linvstd,
                               float& lvar_unb) {
  __shared__ double fin[512];
  constexpr int COLS = 2 * C;
  const int c = threadIdx.x;
  const int lanes = min((int)blockDim.x / COLS, 512 / COLS);
  const int col = c % COLS, lane = c / COLS;
  if (lane < lanes) {
    // deferred mode means at most 64 rows: every load of this thread is issued before the first one is
    // consumed (ONE trip to L2 instead of rows / lanes / 4 of them); the sum keeps its row order
    constexpr int NV = (64 + BLOCK / COLS - 1) / (BLOCK / COLS);
    float v[NV];
    if (LL) {
      const llword* rows = reinterpret_cast<const llword*>(a.part);
      long long t0 = 0;
      for (int spin = 0;; ++spin) {
        llword w[NV];
#pragma unroll
        for (int i = 0; i < NV; ++i) {
          const int r = lane + i * lanes;
          w[i] = r < a.rows ? ll_ld(&rows[(size_t)r * COLS + col]) : ((llword)tag << 32);
        }
        bool ok = true;
#pragma unroll
        for (int i = 0; i < NV; ++i) ok = ok && ((unsigned int)(w[i] >> 32) == tag);
#pragma unroll
        for (int i = 0; i < NV; ++i) v[i] = __uint_as_float((unsigned int)w[i]);
        if (ok) break;
        if (spin == 64) t0 = clock64();
        if (spin > 64 && spin_expired(t0, a.ds)) break;
      }
    } else {
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        const int r = lane + i * lanes;
        v[i] = r < a.rows ? __ldcg(&a.part[(size_t)r * COLS + col]) : 0.f;
      }
    }
    double acc = 0.0;
#pragma unroll
    for (int i = 0; i < NV; ++i) acc += (double)v[i];
    if (!LL)
      for (int r = lane + NV * lanes; r < a.rows; r += lanes) acc += (double)__ldcg(&a.part[(size_t)r * COLS + col]);
    fin[lane * COLS + col] = acc;
  }
  __syncthreads();
  if (c < C) {
    double sm = 0.0, sq = 0.0;
    for (int l = 0; l < lanes; ++l) { sm += fin[l * COLS + 2 * c]; sq += fin[l * COLS + 2 * c + 1]; }
    // sums are combined in double; the few divisions / the square root run in fp32 (an fp64 divide and
    // sqrt chain costs ~1 us of pure latency here, and fp32 is what batch_norm_stats itself uses)
    const double inv_n = (double)(1.0f / local_count);
    const double mean = sm * inv_n;
    double var = sq * inv_n - mean * mean;
    if (var < 0.0) var = 0.0;
    lmean = (float)mean;
    linvstd = 1.0f / sqrtf((float)var + a.eps);
    lvar_unb = (float)var * (local_count / (local_count - 1.0f));
  }
}

// The all_gather of _functions.py:74 as pushed tagged words: this rank's (mean, invstd, unbiased var, count) into
// slot [rank] of every rank's region -- its own included, which is how the other CTAs of this grid get them.
template <int C>
__device__ __forceinline__ void bn_push_stats(const BnPoolArgs& a, int c, float lmean, float linvstd, float lvar_unb,
                                              float local_count, unsigned int tag, bool self_too) {
  for (int q = 0; q < a.comm.world; ++q) {
    if (q == a.comm.rank && !self_too) continue;
    llword* slot = a.comm.peer[q]->ll_bn_fwd[a.layer][a.comm.rank];
    ll_store(&slot[c], lmean, tag);
    ll_store(&slot[C + c], linvstd, tag);
    ll_store(&slot[2 * C + c], lvar_unb, tag);
    if (c == 0) ll_store(&slot[3 * C], local_count, tag);
  }
}

// Small grids, `a.ll`: CTA 0 of the BN kernel calls this BEFORE it waits for the grid dependency.  Under
// programmatic launch this grid is resident while the conv kernel is still finishing, so CTA 0 picks the tagged
// partial rows up as they land, and this rank's statistics are on their way to the peers (one NVLink latency)
// while the kernel boundary is still being crossed.  The other CTAs never touch the rows: after their wait they
// poll the published words of all ranks.  Without the overlap (plain launches) the rows are simply there already.
// What CTA 0 (the CTA that also maintains the running statistics) fetched or computed ahead of the grid dependency.
struct BnOwn {
  bool have = false;                               // this rank's statistics below are valid (bn_forward_publish ran)
  float lmean = 0.f, linvstd = 0.f, lvar_unb = 0.f;
  float rm_old = 0.f, rv_old = 0.f;                // running mean / var of channel threadIdx.x before this step
  long long nbt_old = 0;
};

template <int C, int BLOCK = 256>
__device__ void bn_forward_publish(const BnPoolArgs& a, float local_count, BnOwn& own) {
  const unsigned int tag = (unsigned int)a.ds->epoch + 1u;
  bn_local_stats<C, true, BLOCK>(a, local_count, tag, own.lmean, own.linvstd, own.lvar_unb);
  if (threadIdx.x < C) bn_push_stats<C>(a, threadIdx.x, own.lmean, own.linvstd, own.lvar_unb, local_count, tag, true);
  own.have = true;
}

// CTA 0's early work: the old running statistics (the last kernel that wrote them belongs to the previous step) and,
// with tagged rows, the publication of this rank's statistics.  Must be called by every thread of the CTA.
template <int C, int BLOCK = 256>
__device__ __forceinline__ void bn_forward_early(const BnPoolArgs& a, int rm_off, int rv_off, float local_count, BnOwn& own) {
  if (blockIdx.x != 0 || a.eval) return;
  if (threadIdx.x < C) { own.rm_old = a.bnbuf[rm_off + threadIdx.x]; own.rv_old = a.bnbuf[rv_off + threadIdx.x]; }
  if (threadIdx.x == 0) own.nbt_old = a.nbt[a.layer];
  if (a.ll) bn_forward_publish<C, BLOCK>(a, local_count, own);
}

template <int C>
__device__ void bn_forward_stats(const BnPoolArgs& a, float* s_mean, float* s_invstd, int rm_off, int rv_off,
                                 float local_count, const BnOwn& own) {
  const int c = threadIdx.x;
  const bool xchg = !a.eval && a.sync && a.comm.world > 1;
  const bool defer = !a.eval && a.part != nullptr && !a.ll;
  const bool polled = !a.eval && a.ll;
  float lmean = 0.f, linvstd = 0.f, lvar_unb = 0.f;
  if (defer) {
    bn_local_stats<C, false>(a, local_count, 0u, lmean, linvstd, lvar_unb);
    if (c < C && xchg && blockIdx.x == 0)
      bn_push_stats<C>(a, c, lmean, linvstd, lvar_unb, local_count, (unsigned int)a.ds->epoch + 1u, true);
  } else if (polled) {
    if (own.have) {  // CTA 0 computed them itself
      lmean = own.lmean; linvstd = own.linvstd; lvar_unb = own.lvar_unb;
    } else if (c < C && !xchg) {  // single rank: the words CTA 0 published for this rank
      float got[3];
      const int idx[3] = {c, C + c, 2 * C + c};
      ll_gather<3, 1>(&a.comm.self->ll_bn_fwd[a.layer][a.comm.rank][0], BN_FWD_SLOT, idx, 0, 1,
                      (unsigned int)a.ds->epoch + 1u, got, a.ds);
      lmean = got[0]; linvstd = got[1]; lvar_unb = got[2];
    }
  } else if (!a.eval && c < C) {
    lmean = a.st->lmean[c];
    linvstd = a.st->linvstd[c];
    lvar_unb = a.st->lvar_unbiased[c];
  }
  if (c < C) {
    float mean, invstd, var_unb, n_total = local_count;
    if (a.eval) {
      mean = a.bnbuf[rm_off + c];
      invstd = 1.0f / sqrtf(a.bnbuf[rv_off + c] + a.eps);
      var_unb = 0.f;
    } else if (xchg) {
      float m = 0.f, m2 = 0.f, n = 0.f;
      const unsigned int tag = (unsigned int)a.ds->epoch + 1u;
      float got[MAX_WORLD * 3];
      const int idx[3] = {c, C + c, 3 * C};
      ll_gather<3>(&a.comm.self->ll_bn_fwd[a.layer][0][0], BN_FWD_SLOT, idx, 0, a.comm.world, tag, got, a.ds);
#pragma unroll
      for (int r = 0; r < MAX_WORLD; ++r) {
        if (r >= a.comm.world) break;
        const float mu_r = got[r * 3], is_r = got[r * 3 + 1], n_r = got[r * 3 + 2];
        const float v = (1.0f / (is_r * is_r) - a.eps) * n_r;
        const float f = 1.0f / (n + n_r);
        const float d = m - mu_r;
        m2 = m2 + v + d * d * n * n_r * f;
        m = n * f * m + n_r * f * mu_r;
        n += n_r;
      }
      mean = m;
      invstd = 1.0f / sqrtf(m2 / n + a.eps);
      var_unb = m2 / (n - 1.0f);
      n_total = n;
    } else {
      mean = lmean;
      invstd = linvstd;
      var_unb = lvar_unb;
    }
    s_mean[c] = mean;
    s_invstd[c] = invstd;
    if (blockIdx.x == 0 && !a.eval) {
      a.st->mean[c] = mean;
      a.st->invstd[c] = invstd;
      if (c == 0) a.st->n_total = n_total;
      // a timed-out statistics exchange must not leak into the running buffers (the step is void: k_step)
      if (!xchg || *reinterpret_cast<volatile int*>(&a.ds->error) == 0) {
        if (c == 0) a.nbt[a.layer] = own.nbt_old + 1;
        a.bnbuf[rm_off + c] = (1.0f - a.momentum) * own.rm_old + a.momentum * mean;
        a.bnbuf[rv_off + c] = (1.0f - a.momentum) * own.rv_old + a.momentum * var_unb;
      }
    }
  }
  __syncthreads();
}

template <typename T, int C, int H>
__global__ void __launch_bounds__(256) k_bn_relu_pool(BnPoolArgs a) {
  if (threadIdx.x == 0) TRACE((C == 16 ? 2 : 4) * 16 + 0);
  constexpr int HP = H / 2;
  constexpr int OFF_G = (C == 16) ? OFF_G1 : OFF_G2, OFF_BE = (C == 16) ? OFF_BE1 : OFF_BE2;
  constexpr int RM = (C == 16) ? BNB_RM1 : BNB_RM2, RV = (C == 16) ? BNB_RV1 : BNB_RV2;
  __shared__ float s_mean[C], s_invstd[C], s_g[C], s_b[C];
  if (threadIdx.x < C) {
    s_g[threadIdx.x] = a.params[OFF_G + threadIdx.x];
    s_b[threadIdx.x] = a.params[OFF_BE + threadIdx.x];
  }
  BnOwn own;
  bn_forward_early<C>(a, RM, RV, (float)a.B * (float)(H * H), own);
  pdl_sync();  // predecessor: the conv kernel that produced y and the partial statistics
  const T* __restrict__ y = reinterpret_cast<const T*>(a.y);
  T* __restrict__ p = reinterpret_cast<T*>(a.p);
  const size_t total = (size_t)a.B * C * HP * HP;
  const size_t stride = (size_t)gridDim.x * 256;
  size_t idx0 = (size_t)blockIdx.x * 256 + threadIdx.x;
  // small batch (fewer than four windows per thread): the first window is fetched before the statistics are
  // reduced, so its trip to L2 overlaps the reduction's
  const bool pre = idx0 < total && idx0 + 3 * stride >= total;
  float q0 = 0.f, q1 = 0.f, q2 = 0.f, q3 = 0.f;
  if (pre) {
    const int pw = (int)(idx0 % HP), ph = (int)((idx0 / HP) % HP);
    const T* src = y + ((idx0 / (HP * HP)) * H + 2 * ph) * H + 2 * pw;
    load2<T>(src, q0, q1);
    load2<T>(src + H, q2, q3);
  }
  bn_forward_stats<C>(a, s_mean, s_invstd, RM, RV, (float)a.B * (float)(H * H), own);
  if (threadIdx.x == 0) TRACE((C == 16 ? 2 : 4) * 16 + 1);
  auto emit = [&](size_t idx, float v0, float v1, float v2, float v3) {
    const int pw = (int)(idx % HP), ph = (int)((idx / HP) % HP);
    const size_t plane = idx / (HP * HP);
    const int c = (int)(plane % C);
    const float mu = s_mean[c], is = s_invstd[c], g = s_g[c], b = s_b[c];
    v0 = rnd<T>((v0 - mu) * is * g + b);
    v1 = rnd<T>((v1 - mu) * is * g + b);
    v2 = rnd<T>((v2 - mu) * is * g + b);
    v3 = rnd<T>((v3 - mu) * is * g + b);
    const float m = relu_max4(v0, v1, v2, v3);
    stf<T>(&p[idx], m);
    if (C == 16 && a.p_planes != nullptr) {
      const int n = (int)(plane / C);
      stf<T>(reinterpret_cast<T*>(a.p_planes) + (((size_t)(n * 2 + (c >> 3)) * 320 + (ph + 3) * 16 + pw) * 8 + (c & 7)), m);
    }
  };
  if (pre) { emit(idx0, q0, q1, q2, q3); idx0 += stride; }
  for (; idx0 + 3 * stride < total; idx0 += 4 * stride) {  // large batch: four outputs per iteration, loads first
    float vv[4][4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const size_t idx = idx0 + u * stride;
      const int pw = (int)(idx % HP), ph = (int)((idx / HP) % HP);
      const T* src = y + ((idx / (HP * HP)) * H + 2 * ph) * H + 2 * pw;
      load2<T>(src, vv[u][0], vv[u][1]);
      load2<T>(src + H, vv[u][2], vv[u][3]);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) emit(idx0 + u * stride, vv[u][0], vv[u][1], vv[u][2], vv[u][3]);
  }
  for (; idx0 < total; idx0 += stride) {
    const int pw = (int)(idx0 % HP), ph = (int)((idx0 / HP) % HP);
    const T* src = y + ((idx0 / (HP * HP)) * H + 2 * ph) * H + 2 * pw;
    float v0, v1, v2, v3;
    load2<T>(src, v0, v1);
    load2<T>(src + H, v2, v3);
    emit(idx0, v0, v1, v2, v3);
  }
  if (threadIdx.x == 0) TRACE((C == 16 ? 2 : 4) * 16 + 15);
}

// Same for layer 1 when conv2 runs on the tensor cores: the pooled output is written ONLY in the
// padded-plane layout conv2_tc.cuh consumes ([image][channel/8][20x16 pixels][8 channels]), one
// thread per pixel and group of 8 channels, i.e. one 16-byte store per thread.
template <typename T>
__global__ void __launch_bounds__(256) k_bn1_relu_pool_planes(BnPoolArgs a) {
  constexpr int C = 16, H = 28;
  __shared__ float s_mean[C], s_invstd[C], s_g[C], s_b[C];
  if (threadIdx.x < C) {
    s_g[threadIdx.x] = a.params[OFF_G1 + threadIdx.x];
    s_b[threadIdx.x] = a.params[OFF_BE1 + threadIdx.x];
  }
  BnOwn own;
  bn_forward_early<C>(a, BNB_RM1, BNB_RV1, (float)a.B * (float)(H * H), own);
  pdl_sync();
  bn_forward_stats<C>(a, s_mean, s_invstd, BNB_RM1, BNB_RV1, (float)a.B * (float)(H * H), own);
  const T* __restrict__ y = reinterpret_cast<const T*>(a.y);
  T* __restrict__ pl = reinterpret_cast<T*>(a.p_planes);
  const size_t total = (size_t)a.B * 2 * 196;
  for (size_t idx = (size_t)blockIdx.x * 256 + threadIdx.x; idx < total; idx += (size_t)gridDim.x * 256) {
    const int pw = (int)(idx % 14), ph = (int)((idx / 14) % 14), g = (int)((idx / 196) & 1);
    const size_t n = idx / 392;
    float o[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int c = g * 8 + j;
      const T* src = y + ((n * C + c) * H + 2 * ph) * H + 2 * pw;
      float v0, v1, v2, v3;
      load2<T>(src, v0, v1);
      load2<T>(src + H, v2, v3);
      const float mu = s_mean[c], is = s_invstd[c], gm = s_g[c], b = s_b[c];
      v0 = rnd<T>((v0 - mu) * is * gm + b);
      v1 = rnd<T>((v1 - mu) * is * gm + b);
      v2 = rnd<T>((v2 - mu) * is * gm + b);
      v3 = rnd<T>((v3 - mu) * is * gm + b);
      o[j] = relu_max4(v0, v1, v2, v3);
    }
    T* dst = pl + (((n * 2 + g) * 320 + (ph + 3) * 16 + pw) * 8);
    store4<T>(dst, o[0], o[1], o[2], o[3]);
    store4<T>(dst + 4, o[4], o[5], o[6], o[7]);
  }
}

// ---------------------------------------------------------------------------------------
// 5x5 pad-2 convolution on 14x14 planes, direct form.  MODE 0: conv2 forward (16->32, bias,
// BN statistics).  MODE 2: conv2 data gradient (32->16): the same correlation with the kernel
// flipped and the channel roles swapped.  CTA = one image x 8 output channels; thread = one
// output channel x 2 rows x 7 columns.
struct Conv14Args {
  const void* in;
  const float* params;
  void* out;
  float* part;  // MODE 0: [gridDim.x][32][2]
  int defer;
  BnStat* st;
  DevState* ds;
  CommDev comm;
  int B, dbl, sync;
  float eps;
};

constexpr int C14_RS = 20;  // padded row stride (18 used)

template <typename T, int CI, int CO, int MODE>
__global__ void __launch_bounds__(128) k_conv14(Conv14Args a) {
  extern __shared__ __align__(16) float smem[];
  float* in_s = smem;                      // [CI][18][C14_RS]
  float* w_s = smem + CI * 18 * C14_RS;    // [8][CI][25]
  __shared__ float red[112][2];
  const int tid = threadIdx.x;
  const int cog = blockIdx.y;
  const unsigned int ll_tag = MODE == 0 ? (unsigned int)a.ds->epoch + 1u : 0u;
  const T* __restrict__ in = reinterpret_cast<const T*>(a.in);
  T* __restrict__ out = reinterpret_cast<T*>(a.out);

  for (int i = tid; i < CI * 18 * C14_RS; i += 128) in_s[i] = 0.f;
  for (int i = tid; i < 8 * CI * 25; i += 128) {
    const int k = i % 25, ci = (i / 25) % CI, col = i / (25 * CI);
    float w;
    if (MODE == 0) {
      w = a.params[OFF_W2 + ((cog * 8 + col) * CI + ci) * 25 + k];
    } else {  // out channel (cog*8+col) is conv2's input channel; ci runs over conv2's outputs
      w = a.params[OFF_W2 + (ci * 16 + cog * 8 + col) * 25 + (24 - k)];
    }
    w_s[i] = rnd<T>(w);
  }
  const int col = tid / 14, t = tid % 14;
  const bool active = tid < 112;
  const int rp = t >> 1, c0 = (t & 1) * 7;
  const int co = cog * 8 + col;
  const float bias = (MODE == 0 && active) ? rnd<T>(a.params[OFF_B2 + co]) : 0.f;
  float s = 0.f, ss = 0.f;

  for (int n = blockIdx.x; n < a.B; n += gridDim.x) {
    __syncthreads();
    for (int i = tid; i < CI * 196; i += 128) {
      const int ci = i / 196, r = (i % 196) / 14, c = i % 14;
      in_s[(ci * 18 + r + 2) * C14_RS + c + 2] = ldf<T>(&in[(size_t)n * CI * 196 + i]);
    }
    __syncthreads();
    if (!active) continue;
    float acc[2][7];
#pragma unroll
    for (int r = 0; r < 2; ++r)
#pragma unroll
      for (int c = 0; c < 7; ++c) acc[r][c] = 0.f;
    const float* wt = w_s + col * CI * 25;
#pragma unroll 2
    for (int ci = 0; ci < CI; ++ci) {
      const float* base = in_s + (ci * 18 + 2 * rp) * C14_RS + c0;
#pragma unroll
      for (int ri = 0; ri < 6; ++ri) {
        float row[11];
#pragma unroll
        for (int c = 0; c < 11; ++c) row[c] = base[ri * C14_RS + c];
#pragma unroll
        for (int orow = 0; orow < 2; ++orow) {
          const int kh = ri - orow;
          if (kh >= 0 && kh < 5) {
#pragma unroll
            for (int kw = 0; kw < 5; ++kw) {
              const float w = wt[ci * 25 + kh * 5 + kw];
#pragma unroll
              for (int px = 0; px < 7; ++px) acc[orow][px] = fmaf(row[px + kw], w, acc[orow][px]);
            }
          }
        }
      }
    }
#pragma unroll
    for (int r = 0; r < 2; ++r)
#pragma unroll
      for (int c = 0; c < 7; ++c) {
        float v;
        if (MODE == 0) {
          v = conv_out<T>(acc[r][c], bias, a.dbl);
          s += v;
          ss = fmaf(v, v, ss);
        } else {
          v = acc[r][c];
        }
        stf<T>(&out[(((size_t)n * CO + co) * 14 + 2 * rp + r) * 14 + c0 + c], v);
      }
  }
  if (MODE == 0 && a.part != nullptr) {
    __syncthreads();
    if (active) { red[tid][0] = s; red[tid][1] = ss; }
    __syncthreads();
    if (tid < 16) {
      const int cl = tid >> 1, stat = tid & 1;
      float v = 0.f;
      for (int i = 0; i < 14; ++i) v += red[cl * 14 + i][stat];
      store_partial(a.part, ((size_t)blockIdx.x * 32 + cog * 8 + cl) * 2 + stat, v, a.defer == 2, ll_tag);
    }
    if (!a.defer && last_cta_done(&a.ds->ticket[1], gridDim.x * gridDim.y))
      finalize_fwd_stats<32>(a.part, gridDim.x, (float)a.B * 196.f, a.eps, a.st, 1, a.comm, a.ds, a.sync);
  }
}

// ---------------------------------------------------------------------------------------
// Linear(1568 -> 10) + CrossEntropyLoss (+ the backward of both up to d(pool2 out)).
// CTA = 7 warps, one image at a time; every thread keeps its 7x10 slice of the (rounded)
// fc weight in registers across images.  N = 10 is far too narrow for an MMA tile.
struct FcArgs {
  const void* p2;        // [B,1568]
  const float* params;
  const long long* y;    // labels (may be null in eval)
  const long long* idx;  // null, or this batch's dataset rows: label of image n is y[idx[n]]
  void* logits;          // [B,10] T
  void* dlogits;         // [B,10] T          (train)
  void* dflat;           // [B,1568] T        (train)  d(pool2 out)
  float* nll;            // [B]               (train)
  float* counts;         // [2] correct,size  (eval)
  const DevState* ds;
  int B, use_scaler, eval;
  // Small batch (one image per CTA): BN2's backward pass 1 (`batch_norm_backward_reduce`, _functions.py:144) rides
  // along -- the gradient at the pool output is in this kernel's registers anyway -- and leaves one tagged partial
  // row per image in bpart2 ([32][B][2], see BnBwdArgs in bwd_kernels.cuh).  bpart2 == nullptr: separate kernel.
  const void* y2;       // conv2 output [B,32,14,14]
  const BnStat* st2;    // BN2 statistics of this step
  float* bpart2;
  // ... and, when conv2 left tagged partial rows (`head`), BN2 + ReLU + pool as well: the pooling windows this kernel
  // fetches for the backward pass ARE the seven features each thread feeds to the Linear layer, so the pooled tensor
  // never makes a round trip through L2 and one kernel boundary disappears.  bn2.p receives p2 (the fc weight
  // gradient reads it), CTA 0 plays the BN kernel's CTA 0 (publication, running statistics).
  BnPoolArgs bn2;
  int head;
};

// NI = images per loop iteration: 1 at small batch (shortest chain), 4 at large batch (the two block-wide
// synchronisations per iteration are shared by four images and four times as many loads are in flight).
template <typename T, int NI>
__global__ void __launch_bounds__(224, NI == 4 ? 2 : 1) k_fc_ce(FcArgs a) {  // NI = 4: two CTAs per SM, the 296-CTA grid is one wave
  if (threadIdx.x == 0) TRACE(5 * 16 + 0);
  __shared__ float red[7][NI][10];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const T* __restrict__ p2 = reinterpret_cast<const T*>(a.p2);
  T* __restrict__ logits = reinterpret_cast<T*>(a.logits);
  float w[10][7];
#pragma unroll
  for (int j = 0; j < 10; ++j)
#pragma unroll
    for (int i = 0; i < 7; ++i) w[j][i] = rnd<T>(a.params[OFF_WFC + j * 1568 + tid + 224 * i]);
  float b[10];
#pragma unroll
  for (int j = 0; j < 10; ++j) b[j] = rnd<T>(a.params[OFF_BFC + j]);
  // fused BN2 backward pass 1: this thread's seven pooling windows of conv2's output (two launches old, so they may
  // be fetched before the wait)
  const bool fuse = NI == 1 && a.bpart2 != nullptr && (int)blockIdx.x < a.B;
  const bool head = NI == 1 && a.head != 0;  // implies fuse; grid == B
  float yw[7][4];
  unsigned int ll_tag = 0u;
  auto fetch_windows = [&]() {
    const T* __restrict__ y2 = reinterpret_cast<const T*>(a.y2);
#pragma unroll
    for (int i = 0; i < 7; ++i) {
      const int k = tid + 224 * i, c = k / 49, w = k % 49, ph = w / 7, pw = w % 7;
      const T* src = y2 + (((size_t)blockIdx.x * 32 + c) * 14 + 2 * ph) * 14 + 2 * pw;
      load2<T>(src, yw[i][0], yw[i][1]);
      load2<T>(src + 14, yw[i][2], yw[i][3]);
    }
  };
  BnOwn own;
  if (fuse) ll_tag = (unsigned int)a.ds->epoch + 1u;
  if (head) bn_forward_early<32, 224>(a.bn2, BNB_RM2, BNB_RV2, (float)a.B * 196.f, own);  // CTA 0 publishes BN2's statistics
  else if (fuse) fetch_windows();  // conv2's output is two launches old here
  pdl_sync();  // predecessor: BN2 + ReLU + pool (p2), or conv2 itself (head); the weights above depend on neither
  if (threadIdx.x == 0) TRACE(5 * 16 + 1);
  __shared__ float s_bn[32][4];  // BN2 mean, invstd, gamma, beta
  float pooled[7];
  if (head) {
    __shared__ float s_mean2[32], s_invstd2[32];
    fetch_windows();  // travels together with the poll for the published statistics
    bn_forward_stats<32>(a.bn2, s_mean2, s_invstd2, BNB_RM2, BNB_RV2, (float)a.B * 196.f, own);
    if (tid < 128) {
      const int c = tid >> 2, q = tid & 3;
      s_bn[c][q] = q == 0 ? s_mean2[c] : q == 1 ? s_invstd2[c] : q == 2 ? a.params[OFF_G2 + c] : a.params[OFF_BE2 + c];
    }
    __syncthreads();
    T* __restrict__ p2w = reinterpret_cast<T*>(a.bn2.p);
#pragma unroll
    for (int i = 0; i < 7; ++i) {  // BN2 + ReLU + 2x2 max-pool of this thread's seven windows = its seven fc inputs
      const int k = tid + 224 * i, c = k / 49;
      const float mu = s_bn[c][0], is = s_bn[c][1], gm = s_bn[c][2], bt = s_bn[c][3];
      const float v0 = rnd<T>((yw[i][0] - mu) * is * gm + bt), v1 = rnd<T>((yw[i][1] - mu) * is * gm + bt);
      const float v2 = rnd<T>((yw[i][2] - mu) * is * gm + bt), v3 = rnd<T>((yw[i][3] - mu) * is * gm + bt);
      pooled[i] = relu_max4(v0, v1, v2, v3);
      stf<T>(&p2w[(size_t)blockIdx.x * 1568 + k], pooled[i]);
    }
  } else if (fuse && tid < 128) {  // the first two were written by the predecessor
    const int c = tid >> 2, q = tid & 3;
    s_bn[c][q] = q == 0 ? a.st2->mean[c] : q == 1 ? a.st2->invstd[c] : q == 2 ? a.params[OFF_G2 + c] : a.params[OFF_BE2 + c];
  }
  // incoming gradient of the loss: `scale` cast to the loss dtype (65536 -> inf in fp16)
  const float gl = a.use_scaler ? rnd<T>(a.ds->scale) : 1.0f;
  const float g = rnd<T>(-(gl / (float)a.B));

  for (int n0 = blockIdx.x * NI; n0 < a.B; n0 += gridDim.x * NI) {
    float xk[NI][7];
#pragma unroll
    for (int u = 0; u < NI; ++u)
#pragma unroll
      for (int i = 0; i < 7; ++i)
        xk[u][i] = head ? pooled[i] : (n0 + u < a.B) ? ldf<T>(&p2[(size_t)(n0 + u) * 1568 + tid + 224 * i]) : 0.f;
#pragma unroll
    for (int u = 0; u < NI; ++u)
#pragma unroll
      for (int j = 0; j < 10; ++j) {
        float acc = 0.f;
#pragma unroll
        for (int i = 0; i < 7; ++i) acc = fmaf(xk[u][i], w[j][i], acc);
        acc = warp_sum(acc);
        if (lane == 0) red[warp][u][j] = acc;
      }
    __syncthreads();
#pragma unroll
    for (int u = 0; u < NI; ++u) {
      const int n = n0 + u;
      if (n >= a.B) break;
      float lg[10], mx = -INFINITY;
#pragma unroll
      for (int j = 0; j < 10; ++j) {
        float v = 0.f;
#pragma unroll
        for (int q = 0; q < 7; ++q) v += red[q][u][j];
        lg[j] = rnd<T>(v + b[j]);
        mx = fmaxf(mx, lg[j]);
      }
      if (a.eval) {
#pragma unroll
        for (int j = 0; j < 10; ++j)
          if (tid == j) stf<T>(&logits[(size_t)n * 10 + j], lg[j]);
        if (tid == 0 && a.y != nullptr) {
          // argmax: first maximum, NaN counts as the maximum
          float best = lg[0];
          int arg = 0;
#pragma unroll
          for (int j = 1; j < 10; ++j)
            if (lg[j] > best || (lg[j] != lg[j] && best == best)) { best = lg[j]; arg = j; }
          if ((long long)arg == a.y[a.idx ? a.idx[n] : n]) atomicAdd(&a.counts[0], 1.0f);
          atomicAdd(&a.counts[1], 1.0f);
        }
        continue;
      }
      float se = 0.f;
#pragma unroll
      for (int j = 0; j < 10; ++j) se += expf(lg[j] - mx);
      const float lse = logf(se);
      const int label = (int)a.y[a.idx ? a.idx[n] : n];
      float dl[10], nll = 0.f;
#pragma unroll
      for (int j = 0; j < 10; ++j) {
        const float lsm = rnd<T>(lg[j] - mx - lse);
        if (j == label) nll = -lsm;
        dl[j] = rnd<T>((j == label ? g : 0.f) - expf(lsm) * g);
      }
#pragma unroll
      for (int j = 0; j < 10; ++j)
        if (tid == j) {
          stf<T>(&logits[(size_t)n * 10 + j], lg[j]);
          stf<T>(&reinterpret_cast<T*>(a.dlogits)[(size_t)n * 10 + j], dl[j]);
        }
      if (tid == 0) a.nll[n] = nll;
      T* __restrict__ dflat = reinterpret_cast<T*>(a.dflat);
      float df[7];
#pragma unroll
      for (int i = 0; i < 7; ++i) {
        float acc = 0.f;
#pragma unroll
        for (int j = 0; j < 10; ++j) acc = fmaf(dl[j], w[j][i], acc);
        stf<T>(&dflat[(size_t)n * 1568 + tid + 224 * i], acc);
        df[i] = rnd<T>(acc);
      }
      if (fuse) {
        // per element: the gradient reaches the BN output only through the pool winner, and only if its ReLU is open.
        // Per channel (49 consecutive elements) the two sums are formed in a fixed order: 7 lanes of 7, then the lanes.
        __shared__ float2 s_el[1568];
        __shared__ float2 s_ln[32][7];
#pragma unroll
        for (int i = 0; i < 7; ++i) {
          const int k = tid + 224 * i, c = k / 49;
          const float mu = s_bn[c][0];
          bool live;
          float win;
          pool_winner<T>(yw[i], mu, s_bn[c][1], s_bn[c][2], s_bn[c][3], live, win);
          s_el[k] = live ? make_float2(df[i], df[i] * (win - mu)) : make_float2(0.f, 0.f);
        }
        __syncthreads();
        {
          const int c = tid / 7, l = tid % 7;  // 224 threads = 32 channels x 7 lanes
          float2 acc = make_float2(0.f, 0.f);
#pragma unroll
          for (int q = 0; q < 7; ++q) { const float2 v = s_el[c * 49 + l * 7 + q]; acc.x += v.x; acc.y += v.y; }
          s_ln[c][l] = acc;
        }
        __syncthreads();
        if (tid < 64) {
          const int c = tid >> 1, which = tid & 1;
          float v = 0.f;
#pragma unroll
          for (int l = 0; l < 7; ++l) v += which ? s_ln[c][l].y : s_ln[c][l].x;
          store_partial(a.bpart2, ((size_t)c * a.B + n) * 2 + which, v, 1, ll_tag);
        }
      }
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) TRACE(5 * 16 + 15);
}

}  // namespace dstep
