// conv1 (5x5, pad 2, 1 -> 16 channels on 28x28) as a tcgen05 implicit GEMM for LARGE batches.
//
// On CUDA cores conv1 is FP32-issue bound at large batch (2.6 GFLOP at batch 4096 against ~115 MB of
// traffic: 125 us measured vs 18 us of HBM time).  The tensor cores want a K-major A operand whose
// 16-byte rows are 8 consecutive K elements; with ONE input channel the natural choice is K = (kh, kw):
// build, per image, the table  xe[p] = { R[p], R[p+1], ..., R[p+7] }  (16 bytes, 8 consecutive raster
// pixels in the autocast dtype), where R is the zero-padded 32-wide raster R[32 r + c] = x[r-2][c-2].
// For output pixel m = 32 h + w and filter row kh, the 8 "kw" elements of A's row are exactly xe[m + 32 kh]
// (kw = 5..7 meet zero weights).  Consecutive output pixels = consecutive table entries = the rows of a
// no-swizzle K-major core matrix, so, as in conv2_tc.cuh, every (M-tile, kh pair) is the same table
// addressed through a descriptor with a shifted start address: 7 M-tiles x 3 MMAs (128 x 16 x 16) per image.
//
// Roles (256 threads): warp 0 MMA issuer (+ TMEM owner), warps 1-3 build R and xe for the next image,
// warps 4-7 epilogue (TMEM -> bias/double rounding -> y1 + BN statistics).  Double-buffered table and
// accumulators.  ddp_main.py:17.
#pragma once
#include "conv2_tc.cuh"

namespace dstep {

constexpr int C1T_THREADS = 256;
constexpr int C1T_XE_ENTRIES = 1032;                  // 896 + 4*32 shifts, rounded up
constexpr int C1T_XE_BYTES = C1T_XE_ENTRIES * 16;     // 16,512
constexpr int C1T_R_ELEMS = 1048;                     // raster + 8-element window + slack
constexpr int C1T_W_BYTES = 3 * 512;                  // 3 kh pairs x [2 chunks][16 co][8 kw]
constexpr int C1T_SMEM = 2 * C1T_XE_BYTES + 2 * C1T_R_ELEMS * 2 + C1T_W_BYTES + 256 + 64;

struct Conv1TcArgs {
  BatchSrc x;
  const float* params;
  void* y1;
  float* part;   // [gridDim.x][16][2]
  int defer;
  BnStat* st;
  DevState* ds;
  CommDev comm;
  int B, dbl, sync;
  float eps;
};

template <typename T>
__global__ void __launch_bounds__(C1T_THREADS, 2) k_conv1_tc(Conv1TcArgs a) {
  extern __shared__ __align__(128) unsigned char c1_smem[];
  unsigned char* xe_s = c1_smem;                                   // [2][1032][8] T
  T* r_s = reinterpret_cast<T*>(c1_smem + 2 * C1T_XE_BYTES);       // [2][1048] T
  unsigned char* w_s = c1_smem + 2 * C1T_XE_BYTES + 2 * C1T_R_ELEMS * 2;
  uint64_t* bars = reinterpret_cast<uint64_t*>(w_s + C1T_W_BYTES);
  uint64_t *xfull = bars, *xempty = bars + 2, *tfull = bars + 4, *tempty = bars + 6;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 8);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  // filter as K-major B tiles: [pair][chunk kc][n = co][k8 = kw], kh = 2 pair + kc; zero where kh or kw > 4
  T* wt = reinterpret_cast<T*>(w_s);
  for (int i = tid; i < 3 * 2 * 16 * 8; i += C1T_THREADS) {
    const int k8 = i & 7, n = (i >> 3) & 15, kc = (i >> 7) & 1, pair = i >> 8;
    const int kh = 2 * pair + kc;
    wt[i] = Cvt<T>::from_f((kh < 5 && k8 < 5) ? a.params[OFF_W1 + n * 25 + kh * 5 + k8] : 0.f);
  }
  for (int i = tid; i < 2 * C1T_R_ELEMS; i += C1T_THREADS) r_s[i] = Cvt<T>::from_f(0.f);  // padding stays zero forever
  if (tid == 0) {
    for (int s = 0; s < 2; ++s) {
      mbar_init(&xfull[s], 96);
      mbar_init(&xempty[s], 1);
      mbar_init(&tfull[s], 1);
      mbar_init(&tempty[s], 128);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(256u)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_sync();
  const int n_img = (a.B - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

  if (warp == 0) {
    // ===== MMA issuer =====
    if (lane == 0) {
      const uint32_t idesc = umma_idesc<T>(16);
      const uint64_t bd0 = umma_desc(smem_u32(w_s), 16 * 16, 128);
      for (int i = 0; i < n_img; ++i) {
        const int s = i & 1;
        const uint32_t ph = (i >> 1) & 1;
        mbar_wait(&tempty[s], ph ^ 1);
        mbar_wait(&xfull[s], ph);
        tc_fence_after();
        const uint64_t ad0 = umma_desc(smem_u32(xe_s + s * C1T_XE_BYTES), 32 * 16, 128);
        // third pair = filter row 4 plus a non-existent row 5 (zero weights): point its second K chunk at row 4
        // again (leading offset 0) so it multiplies finite data, not whatever lies past the table
        const uint64_t ad0_last = umma_desc(smem_u32(xe_s + s * C1T_XE_BYTES), 0, 128);
#pragma unroll
        for (int tile = 0; tile < 7; ++tile) {
          const uint32_t d = tmem_base + (uint32_t)((s * 7 + tile) * 16);
#pragma unroll
          for (int pair = 0; pair < 3; ++pair) {
            const uint64_t ad = (pair == 2 ? ad0_last : ad0) + (uint64_t)(((tile * 128 + pair * 64) * 16) >> 4);
            const uint64_t bd = bd0 + (uint64_t)((pair * 512) >> 4);
            umma_f16(d, ad, bd, idesc, pair != 0);
          }
        }
        umma_commit(&xempty[s]);
        umma_commit(&tfull[s]);
      }
    }
  } else if (warp < 4) {
    // ===== builders: image -> zero-padded raster R (autocast dtype) -> table xe =====
    const int bt = tid - 32;  // 0..95
    for (int i = 0; i < n_img; ++i) {
      const int s = i & 1;
      const int n = blockIdx.x + i * gridDim.x;
      T* r = r_s + s * C1T_R_ELEMS;
      mbar_wait(&xempty[s], ((i >> 1) & 1) ^ 1);  // the MMAs that read this table are done (also frees R[s])
      for (int q = bt; q < 196; q += 96) {
        const float4 v = load_pixels4(a.x, n, q);
        const int row = (4 * q) / 28, col = (4 * q) % 28;
        T* d = r + (row + 2) * 32 + col + 2;
        d[0] = Cvt<T>::from_f(v.x); d[1] = Cvt<T>::from_f(v.y); d[2] = Cvt<T>::from_f(v.z); d[3] = Cvt<T>::from_f(v.w);
      }
      asm volatile("bar.sync 1, 96;" ::: "memory");
      const unsigned short* rr = reinterpret_cast<const unsigned short*>(r);
      uint4* xe = reinterpret_cast<uint4*>(xe_s + s * C1T_XE_BYTES);
      for (int p = bt; p < C1T_XE_ENTRIES; p += 96) {
        uint4 o;
        o.x = (uint32_t)rr[p] | ((uint32_t)rr[p + 1] << 16);
        o.y = (uint32_t)rr[p + 2] | ((uint32_t)rr[p + 3] << 16);
        o.z = (uint32_t)rr[p + 4] | ((uint32_t)rr[p + 5] << 16);
        o.w = (uint32_t)rr[p + 6] | ((uint32_t)rr[p + 7] << 16);
        xe[p] = o;
      }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes -> visible to the MMA
      mbar_arrive(&xfull[s]);
      asm volatile("bar.sync 1, 96;" ::: "memory");  // nobody overwrites R[s^1] ... handled by xempty; keeps warps in step
    }
  } else {
    // ===== epilogue =====
    const int q = warp & 3, et = q * 32 + lane;
    T* __restrict__ y1 = reinterpret_cast<T*>(a.y1);
    float bias[16], st32[32];  // st32 = 16 sums, then 16 sums of squares
#pragma unroll
    for (int c = 0; c < 16; ++c) { bias[c] = rnd<T>(a.params[OFF_B1 + c]); st32[c] = 0.f; st32[16 + c] = 0.f; }
    for (int i = 0; i < n_img; ++i) {
      const int s = i & 1;
      const int n = blockIdx.x + i * gridDim.x;
      mbar_wait(&tfull[s], (i >> 1) & 1);
      tc_fence_after();
#pragma unroll 1
      for (int tile = 0; tile < 7; ++tile) {
        float v[16];
        tmem_ld_row<16>(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)((s * 7 + tile) * 16), v);
        const int m = tile * 128 + et, h = m >> 5, w = m & 31;
        if (w < 28) {
#pragma unroll
          for (int c = 0; c < 16; ++c) {
            const float o = conv_out<T>(v[c], bias[c], a.dbl);
            st32[c] += o;
            st32[16 + c] = fmaf(o, o, st32[16 + c]);
            stf<T>(&y1[(((size_t)n * 16 + c) * 28 + h) * 28 + w], o);
          }
        }
      }
      tc_fence_before();
      mbar_arrive(&tempty[s]);
    }
    if (a.part != nullptr) {
      float* red = reinterpret_cast<float*>(xe_s);  // [4 warps][32]; the tables are idle by now
      const float t = warp_transpose_sum32(st32, lane);  // lane l: channel l (sums) / l-16 (squares)
      red[q * 32 + lane] = t;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (a.part != nullptr && tid < 32) {
    const float* red = reinterpret_cast<const float*>(xe_s);
    const float acc = ((red[tid] + red[32 + tid]) + red[64 + tid]) + red[96 + tid];
    const int c = tid & 15, stat = tid >> 4;
    a.part[((size_t)blockIdx.x * 16 + c) * 2 + stat] = acc;
  }
  if (warp == 0) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(256u) : "memory");
  }
  if (a.part != nullptr && !a.defer) {
    if (last_cta_done(&a.ds->ticket[0], gridDim.x))
      finalize_fwd_stats<16>(a.part, gridDim.x, (float)a.B * 784.f, a.eps, a.st, 0, a.comm, a.ds, a.sync);
  }
}

}  // namespace dstep

namespace dstep {

// =======================================================================================
// conv1 weight gradient for LARGE batches on tcgen05, BN1's backward pass 2 fused in front
// (d(conv1 out) never reaches HBM):
//     dW1[co][kh][kw] = sum_{n, q} dy1[n][q][co] * XR[n][q + 30 kh + kw],      db1[co] = sum dy1
// on a 30-wide raster (28 pixels + 2 zero columns that double as the next row's left halo), so the tap
// shift 30 kh + kw stays below 128 and ALL 25 taps fit the M dimension of one accumulator:
//   A (MN-major, M = shift): 16-byte unit j at pixel q = the window table entry xe[q + 8 j]
//                            (8 consecutive raster pixels), units and pixels both 16 bytes apart
//   B (MN-major, N = co 16): d(conv1 out) as two planes [co/8][raster pixel][8 channels]
// D[128 shifts][16 co] lives in TMEM across all images of the CTA: 53 MMAs (128 x 16 x 16) per image, one
// epilogue per CTA.  Warp 0 issues the MMAs; the other 8 warps rebuild, per image, the raster/table from
// x and the dy1 planes from y1, d(pool1 out) and the BN1 statistics (pool/ReLU routing recomputed).
constexpr int C1W_TC_THREADS = 288;
constexpr int C1W_XR_ELEMS = 992;                 // raster (900) + windows
constexpr int C1W_XE_ENTRIES = 976;               // q < 848 plus shifts up to 120, rounded
constexpr int C1W_XE_BYTES = C1W_XE_ENTRIES * 16;
constexpr int C1W_PLANE_PIX = 848;                // 53 K-steps x 16 pixels
constexpr int C1W_PLANE_BYTES = C1W_PLANE_PIX * 16;
constexpr int C1W_DY_BYTES = 2 * C1W_PLANE_BYTES;
constexpr int C1W_TC_SMEM = 2 * (C1W_XE_BYTES + C1W_DY_BYTES + C1W_XR_ELEMS * 2) + 256 + 64;

struct Conv1WgradTcArgs {
  BnBwdArgs bn;  // y = y1, dpool = d(pool1 out), layer = 0, gpart
  BatchSrc x;
};

template <typename T>
__global__ void __launch_bounds__(C1W_TC_THREADS, 2) k_conv1_wgrad_tc(Conv1WgradTcArgs a) {
  extern __shared__ __align__(128) unsigned char c1w_smem[];
  unsigned char* xe_s = c1w_smem;                                        // [2][976][8] T
  unsigned char* dy_s = c1w_smem + 2 * C1W_XE_BYTES;                     // [2][2 planes][848][8] T
  T* r_s = reinterpret_cast<T*>(c1w_smem + 2 * (C1W_XE_BYTES + C1W_DY_BYTES));  // [2][992] T
  uint64_t* bars = reinterpret_cast<uint64_t*>(c1w_smem + 2 * (C1W_XE_BYTES + C1W_DY_BYTES + C1W_XR_ELEMS * 2));
  uint64_t *full = bars, *empty = bars + 2, *done = bars + 4;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 5);
  __shared__ float s_k1[16], s_k2[16];
  __shared__ float s_cst[16][6];  // mean, invstd, gamma, beta, k1, k2
  __shared__ float s_bred[8][16];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  float c_mu = 0.f, c_is = 0.f, c_g = 0.f, c_b = 0.f;
  if (tid < 16) {
    c_mu = a.bn.st->mean[tid]; c_is = a.bn.st->invstd[tid];
    c_g = a.bn.params[OFF_G1 + tid]; c_b = a.bn.params[OFF_BE1 + tid];
  }
  // zero everything once: raster padding, the two pad columns / tail of the dy planes
  for (int i = tid; i < (2 * (C1W_XE_BYTES + C1W_DY_BYTES + C1W_XR_ELEMS * 2)) / 4; i += C1W_TC_THREADS)
    reinterpret_cast<uint32_t*>(c1w_smem)[i] = 0u;
  if (tid == 0) {
    for (int s = 0; s < 2; ++s) { mbar_init(&full[s], 256); mbar_init(&empty[s], 1); }
    mbar_init(done, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(32u)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  bn_backward_sums<16>(a.bn, s_k1, s_k2);  // includes the PDL wait and a __syncthreads
  if (tid < 16) {
    s_cst[tid][0] = c_mu; s_cst[tid][1] = c_is; s_cst[tid][2] = c_g; s_cst[tid][3] = c_b;
    s_cst[tid][4] = s_k1[tid]; s_cst[tid][5] = s_k2[tid];
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int n_img = (a.bn.B - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

  if (warp == 0) {
    if (lane == 0) {
      const uint32_t idesc = umma_idesc_mn<T>(16);
      for (int i = 0; i < n_img; ++i) {
        const int s = i & 1;
        mbar_wait(&full[s], (i >> 1) & 1);
        tc_fence_after();
        const uint64_t ad0 = umma_desc(smem_u32(xe_s + s * C1W_XE_BYTES), 128, 128);
        const uint64_t bd0 = umma_desc(smem_u32(dy_s + s * C1W_DY_BYTES), 128, C1W_PLANE_BYTES);
#pragma unroll
        for (int t = 0; t < 53; ++t)
          umma_f16(tmem_base, ad0 + (uint64_t)((t * 256) >> 4), bd0 + (uint64_t)((t * 256) >> 4), idesc, (i | t) != 0);
        umma_commit(&empty[s]);
      }
      umma_commit(done);
    }
  } else {
    // ===== builders (256 threads) =====
    const int bt = tid - 32;
    const T* __restrict__ y1 = reinterpret_cast<const T*>(a.bn.y);
    const T* __restrict__ dp1 = reinterpret_cast<const T*>(a.bn.dpool);
    float bsum0[8], bsum1[8];  // channel groups 0 and 1
#pragma unroll
    for (int j = 0; j < 8; ++j) { bsum0[j] = 0.f; bsum1[j] = 0.f; }
    for (int i = 0; i < n_img; ++i) {
      const int s = i & 1;
      const int n = blockIdx.x + i * gridDim.x;
      T* r = r_s + s * C1W_XR_ELEMS;
      T* dyp = reinterpret_cast<T*>(dy_s + s * C1W_DY_BYTES);
      mbar_wait(&empty[s], ((i >> 1) & 1) ^ 1);
      if (bt < 196) {  // raster of the input image, autocast dtype
        const float4 v = load_pixels4(a.x, n, bt);
        T* d = r + ((4 * bt) / 28 + 2) * 30 + (4 * bt) % 28 + 2;
        d[0] = Cvt<T>::from_f(v.x); d[1] = Cvt<T>::from_f(v.y); d[2] = Cvt<T>::from_f(v.z); d[3] = Cvt<T>::from_f(v.w);
      }
#pragma unroll
      for (int rep = 0; rep < 2; ++rep) {  // 392 work items: (channel group of 8, pooling window)
        const int item = bt + rep * 256;
        if (item < 392) {
          const int g = item / 196, w = item % 196, ph = w / 14, pw = w % 14;
          float o[4][8];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const int c = g * 8 + j;
            const size_t off = (((size_t)n * 16 + c) * 28 + 2 * ph) * 28 + 2 * pw;
            float v[4];
            load2<T>(y1 + off, v[0], v[1]);
            load2<T>(y1 + off + 28, v[2], v[3]);
            const float dd = ldf<T>(&dp1[((size_t)n * 16 + c) * 196 + w]);
            const float mu = s_cst[c][0], is = s_cst[c][1], gm = s_cst[c][2];
            bool live;
            const int arg = pool_winner<T>(v, mu, is, gm, s_cst[c][3], live);
            const float k1 = s_cst[c][4], k2 = s_cst[c][5], k3 = is * gm;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              const float da = (live && q == arg) ? dd : 0.f;
              o[q][j] = rnd<T>((da - k1 - (v[q] - mu) * k2) * k3);
              if (g == 0) bsum0[j] += o[q][j]; else bsum1[j] += o[q][j];
            }
          }
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            T* dst = dyp + ((size_t)(g * C1W_PLANE_PIX + (2 * ph + (q >> 1)) * 30 + 2 * pw + (q & 1)) * 8);
            store4<T>(dst, o[q][0], o[q][1], o[q][2], o[q][3]);
            store4<T>(dst + 4, o[q][4], o[q][5], o[q][6], o[q][7]);
          }
        }
      }
      asm volatile("bar.sync 1, 256;" ::: "memory");  // raster complete
      const unsigned short* rr = reinterpret_cast<const unsigned short*>(r);
      uint4* xe = reinterpret_cast<uint4*>(xe_s + s * C1W_XE_BYTES);
      for (int p = bt; p < C1W_XE_ENTRIES; p += 256) {
        uint4 o;
        o.x = (uint32_t)rr[p] | ((uint32_t)rr[p + 1] << 16);
        o.y = (uint32_t)rr[p + 2] | ((uint32_t)rr[p + 3] << 16);
        o.z = (uint32_t)rr[p + 4] | ((uint32_t)rr[p + 5] << 16);
        o.w = (uint32_t)rr[p + 6] | ((uint32_t)rr[p + 7] << 16);
        xe[p] = o;
      }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      mbar_arrive(&full[s]);
    }
    // conv1 bias gradient: per-warp transpose-reduce of the 16 channel sums, combined below
    float b32[32];
#pragma unroll
    for (int j = 0; j < 8; ++j) { b32[j] = bsum0[j]; b32[8 + j] = bsum1[j]; b32[16 + j] = 0.f; b32[24 + j] = 0.f; }
    const float t = warp_transpose_sum32(b32, lane);
    if (lane < 16) s_bred[warp - 1][lane] = t;
    // single epilogue: accumulator row = shift 30 kh + kw, column = co
    if (warp <= 4) {
      mbar_wait(done, 0);
      tc_fence_after();
      const int q = warp & 3, sft = q * 32 + lane, kh = sft / 30, kw = sft % 30;
      float v[16];
      tmem_ld_row<16>(tmem_base + ((uint32_t)(q * 32) << 16), v);
      if (kh < 5 && kw < 5) {
        float* row = a.bn.gpart + (size_t)blockIdx.x * N_PAD;
#pragma unroll
        for (int co = 0; co < 16; ++co) row[OFF_W1 + co * 25 + kh * 5 + kw] = v[co];
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (tid < 16) {
    float v = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) v += s_bred[w][tid];
    a.bn.gpart[(size_t)blockIdx.x * N_PAD + OFF_B1 + tid] = v;
  }
  if (warp == 0) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(32u) : "memory");
  }
}

}  // namespace dstep
