// The middle of the backward pass as ONE kernel per (image, M-tile) at small batch:
//     BN2 backward pass 2 (`batch_norm_backward_elemt`, _functions.py:170)  ->  d(conv2 out)
//     conv2 data gradient on tcgen05 (conv2_tc.cuh, MODE 2)                 ->  d(pool1 out)
//     BN1 backward pass 1 (`batch_norm_backward_reduce`, _functions.py:144) ->  per-CTA partial sums
// The three are chained per image -- only the BN2 sums before and the BN1 sums after need the whole batch -- so
// fusing them removes two kernel boundaries and two trips of an activation gradient through L2: d(conv2 out) is
// built directly in the shared-memory planes the MMA reads, d(pool1 out) goes from TMEM through registers into
// the BN1 sums.  d(pool1 out) is still written out for the conv1 weight-gradient kernel; d(conv2 out) is not: the conv2
// weight-gradient branch rebuilds it for itself (k_conv2_wgrad_tc<T, true>) so that it can run beside this kernel.
//
// grid (B, 2): CTA = one image x one 128-row M-tile of the data gradient (8 image rows), 256 threads.
//   all threads   : zero the planes, fetch conv2-output windows (old data: before the grid dependency), then after
//                   the BN2 sums arrived build d(conv2 out) for the image rows the tile's taps touch
//                   (tile 0: pooling rows 0..4, tile 1: 3..6; rows 0..3 / 4..6 are the ones a CTA also writes out)
//   thread 128    : 50 MMAs (25 taps x 2 K-steps of 16 channels), accumulator [128 x 16] in TMEM
//   warps 0..3    : epilogue: TMEM lane = output pixel; d(pool1 out) store + BN1 sums with the pool / ReLU routing
//                   recomputed from conv1's output (fetched while the MMAs run)
// Only used when every partial-row count stays within the tagged-row readers' 64 rows (B <= 32).
#pragma once
#include "bwd_kernels.cuh"
#include "conv2_tc.cuh"

namespace dstep {

struct BwdMidArgs {
  BnBwdArgs bn2;       // y = conv2 out, dpool = d(pool2 out), bpart = BN2 pass-1 rows, dy / dy_planes = d(conv2 out), gpart
  const void* wtiles;  // conv2 filter as data-gradient B tiles (w2_tile_gather)
  void* dp1;           // [B,16,14,14] d(pool1 out)
  const void* y1;      // [B,16,28,28] conv1 out
  const BnStat* st1;   // BN1 statistics of this step
  float* bpart1;       // BN1 pass-1 partial rows, tagged words [16][2B][2]
};

constexpr int MID_THREADS = 256;
constexpr int MID_PLANES_BYTES = 4 * PL_BYTES + TC_SLACK;
constexpr int MID_SMEM_BYTES = TC_WBYTES + MID_PLANES_BYTES + 64;

template <typename T>
__global__ void __launch_bounds__(MID_THREADS, 1) k_bwd_mid_tc(BwdMidArgs a) {
  if (threadIdx.x == 0) TRACE(10 * 16 + 0);
  extern __shared__ __align__(128) unsigned char tc_smem[];
  unsigned char* w_s = tc_smem;
  unsigned char* pl_s = tc_smem + TC_WBYTES;  // d(conv2 out) of this image: [4 planes][320 pixels][8 channels] T
  uint64_t* bars = reinterpret_cast<uint64_t*>(tc_smem + TC_WBYTES + MID_PLANES_BYTES);
  uint64_t *wbar = bars, *tfull = bars + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2);
  __shared__ float s_k1[32], s_k2[32];
  __shared__ float s_bn2[32][4];   // mean, invstd, gamma, beta
  __shared__ float s_bn1[16][4];
  __shared__ float s_ws[32][5][7]; // sum of the four gradients of every window this CTA builds
  __shared__ float s_red[4][32];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int n = blockIdx.x, tile = blockIdx.y, B = a.bn2.B;
  const int ph0 = tile == 0 ? 0 : 3, nph = tile == 0 ? 5 : 4;  // pooling rows whose gradients this tile's taps read
  const int own_lo = tile == 0 ? 0 : 4, own_hi = tile == 0 ? 3 : 6;  // ... and the ones it writes out / sums
  const int nwin = 32 * nph * 7;

  for (int i = tid; i < MID_PLANES_BYTES / 16; i += MID_THREADS) reinterpret_cast<uint4*>(pl_s)[i] = make_uint4(0u, 0u, 0u, 0u);
  if (tid == 0) {
    mbar_init(wbar, 1);
    mbar_init(tfull, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    mbar_expect_tx(wbar, TC_WBYTES);
    tma_bulk_g2s(w_s, a.wtiles, TC_WBYTES, wbar);
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(32u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (tid < 32) { s_bn2[tid][2] = a.bn2.params[OFF_G2 + tid]; s_bn2[tid][3] = a.bn2.params[OFF_BE2 + tid]; }
  if (tid < 16) {  // BN1's statistics are several launches old
    s_bn1[tid][0] = a.st1->mean[tid]; s_bn1[tid][1] = a.st1->invstd[tid];
    s_bn1[tid][2] = a.bn2.params[OFF_G1 + tid]; s_bn1[tid][3] = a.bn2.params[OFF_BE1 + tid];
  }
  const unsigned int ll_tag = (unsigned int)a.bn2.ds->epoch + 1u;
  // this thread's (up to five) pooling windows of conv2's output: produced two or more launches ago
  const T* __restrict__ y2 = reinterpret_cast<const T*>(a.bn2.y);
  const T* __restrict__ dfl = reinterpret_cast<const T*>(a.bn2.dpool);
  float yw[5][4], wd[5];
#pragma unroll
  for (int j = 0; j < 5; ++j) {
    const int wi = min(tid + j * MID_THREADS, nwin - 1);
    const int c = wi / (nph * 7), r = wi % (nph * 7), ph = ph0 + r / 7, pw = r % 7;
    const T* src = y2 + (((size_t)n * 32 + c) * 14 + 2 * ph) * 14 + 2 * pw;
    load2<T>(src, yw[j][0], yw[j][1]);
    load2<T>(src + 14, yw[j][2], yw[j][3]);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // BN2 sums (CTA (0,0) picks the tagged rows up before the wait and publishes); right after the wait: the incoming
  // gradient and BN2's statistics, both possibly written by the direct predecessor
  bn_backward_sums<32>(a.bn2, s_k1, s_k2, [&] {
#pragma unroll
    for (int j = 0; j < 5; ++j) {
      const int wi = min(tid + j * MID_THREADS, nwin - 1);
      const int c = wi / (nph * 7), r = wi % (nph * 7), ph = ph0 + r / 7, pw = r % 7;
      wd[j] = ldf<T>(&dfl[((size_t)n * 32 + c) * 49 + ph * 7 + pw]);
    }
    if (tid < 32) { s_bn2[tid][0] = a.bn2.st->mean[tid]; s_bn2[tid][1] = a.bn2.st->invstd[tid]; }
  });
  if (threadIdx.x == 0) TRACE(10 * 16 + 1);

  // ---- d(conv2 out) into the shared-memory planes (and out to HBM for the weight-gradient kernel)
  T* __restrict__ pl = reinterpret_cast<T*>(pl_s);
  T* __restrict__ dy = reinterpret_cast<T*>(a.bn2.dy);
  T* __restrict__ dyp = reinterpret_cast<T*>(a.bn2.dy_planes);
#pragma unroll
  for (int j = 0; j < 5; ++j) {
    const int wi = tid + j * MID_THREADS;
    if (wi < nwin) {
      const int c = wi / (nph * 7), r = wi % (nph * 7), phl = r / 7, ph = ph0 + phl, pw = r % 7;
      const float mu = s_bn2[c][0], is = s_bn2[c][1], g = s_bn2[c][2], b = s_bn2[c][3];
      bool live;
      const int arg = pool_winner<T>(yw[j], mu, is, g, b, live);
      const float d = live ? wd[j] : 0.f;
      const float k1 = s_k1[c], k2 = s_k2[c], k3 = is * g;
      float o[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float da = (live && i == arg) ? d : 0.f;
        o[i] = rnd<T>((da - k1 - (yw[j][i] - mu) * k2) * k3);
      }
      s_ws[c][phl][pw] = (o[0] + o[1]) + (o[2] + o[3]);
      const int po = (((c >> 3) * PL_PIX) + (2 * ph + 3) * 16 + 2 * pw) * 8 + (c & 7);
      stf<T>(pl + po, o[0]); stf<T>(pl + po + 8, o[1]); stf<T>(pl + po + 128, o[2]); stf<T>(pl + po + 136, o[3]);
      if (dy != nullptr && ph >= own_lo && ph <= own_hi) {  // only when somebody downstream reads it from HBM
        const size_t off = (((size_t)n * 32 + c) * 14 + 2 * ph) * 14 + 2 * pw;
        store2<T>(dy + off, o[0], o[1]);
        store2<T>(dy + off + 14, o[2], o[3]);
        T* gp = dyp + (size_t)n * 4 * PL_PIX * 8 + po;
        stf<T>(gp, o[0]); stf<T>(gp + 8, o[1]); stf<T>(gp + 128, o[2]); stf<T>(gp + 136, o[3]);
      }
    }
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes of the planes -> visible to UMMA
  __syncthreads();
  if (threadIdx.x == 0) TRACE(10 * 16 + 2);

  // conv2 bias gradient partial of this CTA: its own windows, fixed order
  if (tid < 32) {
    float v = 0.f;
    for (int ph = own_lo; ph <= own_hi; ++ph)
#pragma unroll
      for (int pw = 0; pw < 7; ++pw) v += s_ws[tid][ph - ph0][pw];
    a.bn2.gpart[(size_t)(tile * B + n) * N_PAD + OFF_B2 + tid] = v;
  }

  if (tid == 128) {
    // ===== MMA issuer: 25 taps x 2 K-steps, every descriptor a compile-time offset from the plane base =====
    const uint32_t idesc = umma_idesc<T>(16);
    mbar_wait(wbar, 0);
    tc_fence_after();
    const uint64_t ad0 = umma_desc(smem_u32(pl_s), PL_BYTES, 128);
    const uint64_t bd0 = umma_desc(smem_u32(w_s), 16 * 16, 128);
    const uint64_t tile_off = (uint64_t)((tile * 128 * 16) >> 4);
#pragma unroll
    for (int tap = 0; tap < 25; ++tap) {
      const int kh = tap / 5, kw = tap % 5;
      const uint32_t shift = (uint32_t)(16 * (kh + 1) + kw - 2) * 16u;
#pragma unroll
      for (int kk = 0; kk < 2; ++kk) {
        const uint64_t ad = ad0 + tile_off + (uint64_t)((shift + kk * 2 * PL_BYTES) >> 4);
        const uint64_t bd = bd0 + (uint64_t)((tap * 1024 + kk * 2 * 16 * 16) >> 4);
        umma_f16(tmem_base, ad, bd, idesc, (tap | kk) != 0);
      }
    }
    umma_commit(tfull);
  }

  if (warp < 4) {
    // ===== epilogue: one output pixel per thread =====
    const int m = tile * 128 + warp * 32 + lane, h = m >> 4, w = m & 15;
    const bool ok = h < 14 && w < 14;
    const T* __restrict__ y1 = reinterpret_cast<const T*>(a.y1);
    float yv[16][4];
#pragma unroll
    for (int c = 0; c < 16; ++c) {  // conv1's output at the 16 pooling windows this pixel is the output of
      const T* src = y1 + (((size_t)n * 16 + c) * 28 + 2 * (ok ? h : 0)) * 28 + 2 * (ok ? w : 0);
      load2<T>(src, yv[c][0], yv[c][1]);
      load2<T>(src + 28, yv[c][2], yv[c][3]);
    }
    mbar_wait(tfull, 0);
    tc_fence_after();
    float v[16];
    tmem_ld_row<16>(tmem_base + ((uint32_t)(warp * 32) << 16), v);
    T* __restrict__ dp1 = reinterpret_cast<T*>(a.dp1);
    float sums[32];
#pragma unroll
    for (int c = 0; c < 16; ++c) {
      const float d = rnd<T>(v[c]);
      bool live;
      float win;
      pool_winner<T>(yv[c], s_bn1[c][0], s_bn1[c][1], s_bn1[c][2], s_bn1[c][3], live, win);
      const bool use = ok && live;
      sums[c] = use ? d : 0.f;
      sums[16 + c] = use ? d * (win - s_bn1[c][0]) : 0.f;
      if (ok) stf<T>(&dp1[(((size_t)n * 16 + c) * 14 + h) * 14 + w], v[c]);
    }
    const float t = warp_transpose_sum32(sums, lane);  // lane l: sum over the warp's 32 pixels of entry l
    s_red[warp][lane] = t;
  }
  tc_fence_before();
  __syncthreads();
  if (tid < 32) {  // entry l < 16: sum dy of channel l; l >= 16: sum dy * (x - mean) of channel l - 16
    const float acc = ((s_red[0][tid] + s_red[1][tid]) + s_red[2][tid]) + s_red[3][tid];
    const int c = tid & 15, which = tid >> 4;
    store_partial(a.bpart1, ((size_t)c * (2 * B) + (tile * B + n)) * 2 + which, acc, 1, ll_tag);
  }
  if (warp == 0) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(32u) : "memory");
    if (threadIdx.x == 0) TRACE(10 * 16 + 15);
  }
}

}  // namespace dstep
