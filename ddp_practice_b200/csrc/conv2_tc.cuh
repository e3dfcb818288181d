// conv2 (5x5, pad 2, 16<->32 channels on 14x14 planes) as an implicit GEMM on the 5th-gen
// tensor cores: tcgen05.mma with the accumulator in TMEM, operands in shared memory, the
// image brought in by one TMA bulk copy, a three-role warp-specialised persistent CTA.
//
// No im2col is ever materialised.  Activations live in HBM in a "padded plane" layout
//     [image][channel/8][20 rows x 16 pixels][8 channels]      (16 B per pixel and plane)
// with image row h at plane row h+3, columns 14..15 and rows 0..2 / 17..19 zero.  Rows being
// 16 pixels wide, the two zero columns serve as right halo of one row AND left halo of the
// next, so for output pixel m = h*16 + w the input pixel of filter tap (kh,kw) is simply
// plane pixel  m + 16*(kh+1) + kw - 2 : a constant shift.  In the K-major, no-swizzle UMMA
// shared-memory layout one core matrix is 8 rows x 16 B stored contiguously -- exactly 8
// consecutive pixels of one plane -- so the A operand of every tap is THE SAME staged image
// addressed through a descriptor whose start address is moved by shift*16 B.  One 128-row MMA
// covers 8 image rows; an image is two M-tiles (rows 224..255 and columns 14..15 are discarded
// in the epilogue).  Per image: 2 tiles x 25 taps x (Cin/16) MMAs of 128 x N x 16.
//
//   MODE 0  forward        A = p1 planes (Cin 16, 2 planes)   N = 32   D -> y2  (+ BN statistics)
//   MODE 2  data gradient  A = dy2 planes (Cin 32, 4 planes)  N = 16   D -> d(pool1 out)
//
// ddp_main.py:23 (forward) and its autograd (ddp_main.py:91).
#pragma once
#include "bwd_kernels.cuh"

namespace dstep {

constexpr int PL_ROWS = 20, PL_PIX = PL_ROWS * 16;     // 320 pixels per plane
constexpr int PL_BYTES = PL_PIX * 16;                  // 5120 B per plane
constexpr int TC_THREADS = 192;                        // warp 0 TMA, warp 1 MMA, warps 2..5 epilogue
constexpr int TC_WBYTES = 25 * 1024;                   // 25 taps x (N x K x 2 B = 1 KB)
constexpr int TC_SLACK = 512;                          // tile 1 of the last plane reads 18 pixels past the plane

__host__ __device__ constexpr int tc_stage_bytes(int planes) { return planes * PL_BYTES + TC_SLACK; }
__host__ __device__ constexpr int tc_smem_bytes(int planes) { return TC_WBYTES + 2 * tc_stage_bytes(planes) + 256; }

// element offset of (image n, channel c, row h, col w) in a padded-plane tensor with `planes` planes
__device__ __forceinline__ size_t plane_off(int n, int planes, int c, int h, int w) {
  return ((size_t)(n * planes + (c >> 3)) * PL_PIX + (h + 3) * 16 + w) * 8 + (c & 7);
}

// ---- PTX wrappers ---------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc),
      "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {  // arrives when all MMAs issued so far are done
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// shared-memory matrix descriptor, K-major, no swizzle (cute::UMMA::SmemDescriptor):
// bits [0,14) start>>4, [16,30) leading byte offset>>4 (between the two 8-element K chunks),
// [32,46) stride byte offset>>4 (between 8-row groups), [46,48) version = 1, [61,64) layout 0
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16) |
         ((uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32) | (1ull << 46);
}
// instruction descriptor (cute::UMMA::InstrDescriptor): D fp32, A/B fp16 or bf16, both K-major, M = 128
template <typename T> __device__ __forceinline__ uint32_t umma_idesc(int n);
template <> __device__ __forceinline__ uint32_t umma_idesc<__half>(int n) {
  return (1u << 4) | (0u << 7) | (0u << 10) | ((uint32_t)(n >> 3) << 17) | ((128u >> 4) << 24);
}
template <> __device__ __forceinline__ uint32_t umma_idesc<__nv_bfloat16>(int n) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((128u >> 4) << 24);
}

template <int N> __device__ __forceinline__ void tmem_ld_row(uint32_t taddr, float* v);
template <> __device__ __forceinline__ void tmem_ld_row<16>(uint32_t taddr, float* v) {
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}
template <> __device__ __forceinline__ void tmem_ld_row<32>(uint32_t taddr, float* v) {
  tmem_ld_row<16>(taddr, v);
  tmem_ld_row<16>(taddr + 16, v + 16);
}

// Sum v[c] over the 32 lanes of a warp for all 32 channels c at once; lane l returns channel l's
// total.  31 shuffles instead of 32 x 5: at every step a lane keeps the half of its values whose
// index bit matches its lane bit and trades the other half with its partner.  Fixed order.
__device__ __forceinline__ float warp_transpose_sum32(float* v, int lane) {
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) {
    const bool up = (lane & off) != 0;
#pragma unroll
    for (int i = 0; i < off; ++i) {
      const float keep = up ? v[i + off] : v[i];
      const float send = up ? v[i] : v[i + off];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
    }
  }
  return v[0];
}

struct Conv2TcArgs {
  const void* wtiles;  // [25][1 KB] B operand tiles for this MODE (step_kernel.cuh: w2_tile_store)
  const void* in;      // padded planes: MODE 0 p1h [B][2][320][8], MODE 2 dy2h [B][4][320][8]
  const float* params;
  void* out;           // MODE 0 y2 [B,32,14,14], MODE 2 dp1 [B,16,14,14]  (NCHW, T)
  float* part;         // MODE 0: BN partial sums [gridDim.x][32][2] (nullptr in eval)
  int defer;           // 1: the consumer reduces the partial rows (no last-CTA tail here)
  BnStat* st;
  DevState* ds;
  CommDev comm;
  int B, dbl, sync;
  float eps;
};

template <typename T, int MODE>
__global__ void __launch_bounds__(TC_THREADS, 2) k_conv2_tc(Conv2TcArgs a) {
  if (threadIdx.x == 0) TRACE((MODE == 0 ? 3 : 10) * 16 + 0);
  constexpr int PLANES = MODE == 0 ? 2 : 4;   // input channel groups of 8
  constexpr int N = MODE == 0 ? 32 : 16;      // output channels
  constexpr int KSTEPS = PLANES / 2;          // MMAs (K = 16) per tap
  constexpr int STAGE = tc_stage_bytes(PLANES);
  constexpr uint32_t IMG_BYTES = PLANES * PL_BYTES;
  extern __shared__ __align__(128) unsigned char tc_smem[];
  unsigned char* w_s = tc_smem;                         // [25][KSTEPS*2 chunks][N rows][8] T
  unsigned char* in_s = tc_smem + TC_WBYTES;            // [2 stages][PLANES][320][8] T (+ slack)
  uint64_t* bars = reinterpret_cast<uint64_t*>(tc_smem + TC_WBYTES + 2 * STAGE);
  uint64_t *full = bars, *empty = bars + 2, *tfull = bars + 4, *tempty = bars + 6, *wbar = bars + 8;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 9);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  // ---- one-time setup: barriers, TMEM, and the 25 filter taps (ready-made K-major B tiles kept
  // up to date by the SGD kernel) fetched with one TMA copy
  // the slack after each stage is read by discarded rows only, but must not hold NaN patterns that
  // could trap nothing -- it is plain data; zero it once for reproducible garbage rows
  for (int i = tid; i < 2 * (TC_SLACK / 4); i += TC_THREADS) {
    const int s = i / (TC_SLACK / 4), j = i % (TC_SLACK / 4);
    reinterpret_cast<uint32_t*>(in_s + s * STAGE + IMG_BYTES)[j] = 0u;
  }
  if (tid == 0) {
    for (int s = 0; s < 2; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
      mbar_init(&tfull[s], 1);
      mbar_init(&tempty[s], 128);
    }
    mbar_init(wbar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    mbar_expect_tx(wbar, TC_WBYTES);
    tma_bulk_g2s(w_s, a.wtiles, TC_WBYTES, wbar);
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(128u)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes of w_s -> visible to UMMA
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const unsigned int ll_tag = (unsigned int)a.ds->epoch + 1u;  // for the partial-row stores at the very end
  pdl_sync();  // predecessor wrote the input planes; barriers, TMEM and the filter tiles did not need it
  if (threadIdx.x == 0) TRACE((MODE == 0 ? 3 : 10) * 16 + 1);
  const int n_img = (a.B - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;  // images of this CTA
  // gridDim.y == 2 (small batches): each CTA computes ONE of the two M-tiles of its images, so twice as many SMs
  // share the MMA and epilogue work; gridDim.y == 1: both tiles, every image is fetched once
  const int tile_lo = gridDim.y == 2 ? (int)blockIdx.y : 0, tile_hi = gridDim.y == 2 ? (int)blockIdx.y + 1 : 2;
  const unsigned char* gin = reinterpret_cast<const unsigned char*>(a.in);

  if (warp == 0) {
    // ===== TMA producer: one bulk copy per image =====
    if (lane == 0) {
      for (int i = 0; i < n_img; ++i) {
        const int s = i & 1;
        mbar_wait(&empty[s], ((i >> 1) & 1) ^ 1);
        mbar_expect_tx(&full[s], IMG_BYTES);
        const int n = blockIdx.x + i * gridDim.x;
        tma_bulk_g2s(in_s + s * STAGE, gin + (size_t)n * IMG_BYTES, IMG_BYTES, &full[s]);
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: 2 tiles x 25 taps x KSTEPS instructions per image =====
    if (lane == 0) {
      const uint32_t idesc = umma_idesc<T>(N);
      const uint32_t w_addr = smem_u32(w_s);
      mbar_wait(wbar, 0);
      for (int i = 0; i < n_img; ++i) {
        const int s = i & 1;
        const uint32_t ph = (i >> 1) & 1;
        mbar_wait(&tempty[s], ph ^ 1);
        mbar_wait(&full[s], ph);
        tc_fence_after();
        // descriptors differ from a per-image base only by compile-time constants: the whole tap loop is
        // unrolled so each MMA costs one 64-bit add per operand (a div/mod per tap made the single issuing
        // thread, not the tensor core, the bottleneck: measured ~160 cycles per MMA)
        const uint64_t ad0 = umma_desc(smem_u32(in_s + s * STAGE), PL_BYTES, 128);
        const uint64_t bd0 = umma_desc(w_addr, N * 16, 128);
#pragma unroll
        for (int tile = 0; tile < 2; ++tile) {
          if (tile < tile_lo || tile >= tile_hi) continue;
          const uint32_t d = tmem_base + (uint32_t)((s * 2 + tile) * N);
#pragma unroll
          for (int tap = 0; tap < 25; ++tap) {
            const int kh = tap / 5, kw = tap % 5;
            const uint32_t shift = (uint32_t)(tile * 128 + 16 * (kh + 1) + kw - 2) * 16u;
#pragma unroll
            for (int kk = 0; kk < KSTEPS; ++kk) {
              const uint64_t ad = ad0 + (uint64_t)((shift + kk * 2 * PL_BYTES) >> 4);
              const uint64_t bd = bd0 + (uint64_t)((tap * 1024 + kk * 2 * N * 16) >> 4);
              umma_f16(d, ad, bd, idesc, (tap | kk) != 0);
            }
          }
        }
        umma_commit(&empty[s]);   // smem stage may be refilled
        umma_commit(&tfull[s]);   // accumulators ready
      }
    }
  } else {
    // ===== epilogue: TMEM -> registers -> bias / rounding / BN statistics -> HBM =====
    const int q = warp & 3;                      // TMEM lane quarter this warp may touch
    const int et = q * 32 + lane;                // row within the tile
    T* __restrict__ out = reinterpret_cast<T*>(a.out);
    float bias[MODE == 0 ? 32 : 1];
    float s_sum[MODE == 0 ? 32 : 1], s_sq[MODE == 0 ? 32 : 1];
    if (MODE == 0) {
#pragma unroll
      for (int c = 0; c < 32; ++c) { bias[c] = rnd<T>(a.params[OFF_B2 + c]); s_sum[c] = 0.f; s_sq[c] = 0.f; }
    }
    for (int i = 0; i < n_img; ++i) {
      const int s = i & 1;
      const int n = blockIdx.x + i * gridDim.x;
      mbar_wait(&tfull[s], (i >> 1) & 1);
      tc_fence_after();
#pragma unroll 1
      for (int tile = tile_lo; tile < tile_hi; ++tile) {
        float v[N];
        tmem_ld_row<N>(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)((s * 2 + tile) * N), v);
        const int m = tile * 128 + et, h = m >> 4, w = m & 15;
        if (h < 14 && w < 14) {
#pragma unroll
          for (int c = 0; c < N; ++c) {
            float o;
            if (MODE == 0) {
              o = conv_out<T>(v[c], bias[c], a.dbl);
              s_sum[c] += o;
              s_sq[c] = fmaf(o, o, s_sq[c]);
            } else {
              o = v[c];
            }
            stf<T>(&out[(((size_t)n * N + c) * 14 + h) * 14 + w], o);
          }
        }
      }
      tc_fence_before();
      mbar_arrive(&tempty[s]);
    }
    if (MODE == 0 && a.part != nullptr) {
      // deterministic CTA reduction of the per-thread statistics: butterfly inside the warp, then
      // the four epilogue warps are summed in a fixed order below
      float* red = reinterpret_cast<float*>(in_s);  // [4 warps][64]; the image stages are idle by now
      const float t0 = warp_transpose_sum32(s_sum, lane), t1 = warp_transpose_sum32(s_sq, lane);
      red[q * 64 + lane] = t0;        // lane l ends up with channel l
      red[q * 64 + 32 + lane] = t1;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (MODE == 0 && a.part != nullptr) {
    const float* red = reinterpret_cast<const float*>(in_s);
    if (tid < 64) {
      const float acc = ((red[tid] + red[64 + tid]) + red[128 + tid]) + red[192 + tid];
      const int c = tid & 31, stat = tid >> 5;
      store_partial(a.part, ((size_t)(blockIdx.y * gridDim.x + blockIdx.x) * 32 + c) * 2 + stat, acc, a.defer == 2, ll_tag);
    }
  }
  if (warp == 0) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(128u) : "memory");
    if (threadIdx.x == 0) TRACE((MODE == 0 ? 3 : 10) * 16 + 15);
  }
  if (MODE == 0 && a.part != nullptr) {
    if (!a.defer && last_cta_done(&a.ds->ticket[1], gridDim.x * gridDim.y))
      finalize_fwd_stats<32>(a.part, gridDim.x * gridDim.y, (float)a.B * 196.f, a.eps, a.st, 1, a.comm, a.ds, a.sync);
  }
}


// =======================================================================================
// conv2 weight gradient on tcgen05:
//     dW2[co][ci][kh][kw] = sum_{n, pixels q} dy2[n][q][co] * p1[n][q + (kh-2)*16 + (kw-2)][ci]
// The reduction runs over pixels, so both operands are used MN-major (the 16-byte unit of the
// padded-plane layout = 8 channels of one pixel = 8 consecutive M resp. N elements; consecutive
// pixels = consecutive K):
//   A = p1 plane c8  M = (shift j, ci%8): M-unit j is the SAME plane moved by j pixels -- a stride of
//                    one 16-byte unit between M-units, i.e. overlapping core matrices.  M = 128 gives
//                    16 shifts; j = 0..4 are the five kw taps, the other accumulator rows are never read.
//   B = dy2 planes   N = co = 32 (4 units, one per plane)
// One accumulator [128 x 32] per (kh, c8), kept in TMEM across ALL images of the CTA; 14 K-steps of
// 16 pixels cover the 224-pixel output raster (its two pad columns hold zeros in dy2).  140 MMAs per
// image; one epilogue per CTA writes one partial row.  gridDim.y = 5 splits the kh taps over 5 CTAs
// (small batches: more SMs, 64 TMEM columns each); gridDim.y = 1 keeps all ten accumulators (320
// columns) in one CTA (large batches: every image is fetched once).
struct Conv2WgradTcArgs {
  const void* dy2h;  // [B][4][320][8]
  const void* p1h;   // [B][2][320][8]
  float* gpart;      // [rows = gridDim.x][N_PAD]
  int B;
  // BUILD = true (small batch, together with the fused backward middle, bwd_mid_tc.cuh): d(conv2 out) is not read from
  // HBM but rebuilt here from conv2's output, the gradient at the pool output and the BN2 backward sums (the same
  // `batch_norm_backward_elemt` arithmetic as k_bn2_bwd_apply), so this branch of the step graph depends on the
  // Linear+CE kernel only and runs BESIDE the backward middle instead of behind it.
  BnBwdArgs bn2;     // y = conv2 out, dpool = d(pool2 out), st = BN2 statistics, defer = 2 (sums are polled)
};

constexpr int WG_DY_STAGE = 4 * PL_BYTES, WG_P_STAGE = 2 * PL_BYTES;
constexpr int WG_BAR_OFF = 2 * WG_DY_STAGE + 2 * WG_P_STAGE + 1024;   // 1 KB slack: shifts read past the last p plane
constexpr int WG_SMEM_BYTES = WG_BAR_OFF + 256;

// MN-major operands: bits 15 / 16 of the instruction descriptor; in the shared-memory descriptor
// LBO = stride between groups of 8 K, SBO = stride between 16-byte units along M/N
// (cute::UMMA::make_umma_desc<Major::MN>)
template <typename T> __device__ __forceinline__ uint32_t umma_idesc_mn(int n) {
  return umma_idesc<T>(n) | (1u << 15) | (1u << 16);
}

constexpr int WG_BUILD_THREADS = 64 + 320;  // BUILD: warp 0 TMA, warp 1 MMA, warps 2..11 builders (warps 4, 5 also the epilogue)

template <typename T, bool BUILD>
__global__ void __launch_bounds__(BUILD ? WG_BUILD_THREADS : TC_THREADS, BUILD ? 2 : 1) k_conv2_wgrad_tc(Conv2WgradTcArgs a) {
  constexpr int NTHREADS = BUILD ? WG_BUILD_THREADS : TC_THREADS;
  constexpr int NBUILD = NTHREADS - 64;
  if (threadIdx.x == 0) TRACE(8 * 16 + 0);
  __shared__ float s_k1[32], s_k2[32];
  __shared__ float s_bn2[32][4];  // BN2 mean, invstd, gamma, beta (BUILD)
  extern __shared__ __align__(128) unsigned char wg_smem[];
  unsigned char* dy_s = wg_smem;                         // [2][4 planes]
  unsigned char* p_s = wg_smem + 2 * WG_DY_STAGE;        // [2][2 planes] + slack
  uint64_t* bars = reinterpret_cast<uint64_t*>(wg_smem + WG_BAR_OFF);
  uint64_t *full = bars, *empty = bars + 2, *done = bars + 4;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 5);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int kh_lo = gridDim.y == 1 ? 0 : (int)blockIdx.y, kh_n = gridDim.y == 1 ? 5 : 1;
  const uint32_t tmem_cols = gridDim.y == 1 ? 512u : 64u;
  for (int i = tid; i < 256; i += NTHREADS) reinterpret_cast<uint32_t*>(p_s + 2 * WG_P_STAGE)[i] = 0u;
  if (BUILD)  // the builders only ever write the 14 x 14 interior: halo rows / columns of both stages stay zero
    for (int i = tid; i < 2 * WG_DY_STAGE / 16; i += NTHREADS) reinterpret_cast<uint4*>(dy_s)[i] = make_uint4(0u, 0u, 0u, 0u);
  if (tid == 0) {
    // BUILD: a stage is full when the p1 planes have landed (TMA) AND the four builder warps have written d(conv2 out)
    for (int s = 0; s < 2; ++s) { mbar_init(&full[s], BUILD ? 1 + NBUILD / 32 : 1); mbar_init(&empty[s], 1); }
    mbar_init(done, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(tmem_cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int n_img = (a.B - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
  const unsigned char* gdy = reinterpret_cast<const unsigned char*>(a.dy2h);
  const unsigned char* gp = reinterpret_cast<const unsigned char*>(a.p1h);

  // BUILD: this thread's (up to 13) pooling windows of one image: conv2's output and the gradient at the pool output
  constexpr int NWIN = (32 * 49 + NBUILD - 1) / NBUILD;  // windows per builder thread (5 with 320 builders)
  const int bt = tid - 64;  // builder thread index (warps 2 and up)
  float yw[BUILD ? NWIN : 1][4], wd[BUILD ? NWIN : 1];
  auto fetch_windows = [&](int n) {
    if constexpr (BUILD) {
      const T* __restrict__ y2 = reinterpret_cast<const T*>(a.bn2.y);
      const T* __restrict__ dfl = reinterpret_cast<const T*>(a.bn2.dpool);
#pragma unroll
      for (int j = 0; j < NWIN; ++j) {
        const int wi = min(bt + j * NBUILD, 32 * 49 - 1);
        const int c = wi / 49, r = wi % 49, ph = r / 7, pw = r % 7;
        const T* src = y2 + (((size_t)n * 32 + c) * 14 + 2 * ph) * 14 + 2 * pw;
        load2<T>(src, yw[j][0], yw[j][1]);
        load2<T>(src + 14, yw[j][2], yw[j][3]);
        wd[j] = ldf<T>(&dfl[((size_t)n * 32 + c) * 49 + r]);
      }
    }
  };
  if constexpr (BUILD) {
    // everything this kernel reads was written before its (full) graph dependency, the Linear+CE kernel -- except the
    // BN2 backward sums, which CTA (0,0) of the backward-middle kernel running beside it publishes as tagged words
    if (warp >= 2 && n_img > 0) fetch_windows(blockIdx.x);
    bn_backward_sums<32>(a.bn2, s_k1, s_k2, [] {}, true);
    if (tid < 32) {
      s_bn2[tid][0] = a.bn2.st->mean[tid]; s_bn2[tid][1] = a.bn2.st->invstd[tid];
      s_bn2[tid][2] = a.bn2.params[OFF_G2 + tid]; s_bn2[tid][3] = a.bn2.params[OFF_BE2 + tid];
    }
    __syncthreads();
    if (threadIdx.x == 0) TRACE(8 * 16 + 1);
  }

  if (warp == 0) {
    if (lane == 0) {
      for (int i = 0; i < n_img; ++i) {
        const int s = i & 1;
        mbar_wait(&empty[s], ((i >> 1) & 1) ^ 1);
        mbar_expect_tx(&full[s], (BUILD ? 0 : WG_DY_STAGE) + WG_P_STAGE);
        const int n = blockIdx.x + i * gridDim.x;
        if (!BUILD) tma_bulk_g2s(dy_s + s * WG_DY_STAGE, gdy + (size_t)n * WG_DY_STAGE, WG_DY_STAGE, &full[s]);
        tma_bulk_g2s(p_s + s * WG_P_STAGE, gp + (size_t)n * WG_P_STAGE, WG_P_STAGE, &full[s]);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc = umma_idesc_mn<T>(32);
      for (int i = 0; i < n_img; ++i) {
        const int s = i & 1;
        mbar_wait(&full[s], (i >> 1) & 1);
        tc_fence_after();
        const uint64_t bd0 = umma_desc(smem_u32(dy_s + s * WG_DY_STAGE) + 48 * 16, 128, PL_BYTES);
#pragma unroll 1
        for (int k = 0; k < kh_n; ++k) {
          const int kh = kh_lo + k;
          // first output pixel of K-step t is plane raster index 48 + 16 t; the p1 operand starts (kh-2) rows
          // and 2 pixels earlier; everything else is a compile-time constant added to the descriptor
          const uint64_t ad0 = umma_desc(smem_u32(p_s + s * WG_P_STAGE) + (48 + (kh - 2) * 16 - 2) * 16, 128, 16);
#pragma unroll
          for (int c8 = 0; c8 < 2; ++c8) {
            const uint32_t d = tmem_base + (uint32_t)((k * 2 + c8) * 32);
#pragma unroll
            for (int t = 0; t < 14; ++t) {
              const uint64_t ad = ad0 + (uint64_t)((c8 * PL_BYTES + t * 256) >> 4);
              const uint64_t bd = bd0 + (uint64_t)((t * 256) >> 4);
              umma_f16(d, ad, bd, idesc, (i | t) != 0);
            }
          }
        }
        umma_commit(&empty[s]);
      }
      umma_commit(done);
    }
  }
  if (BUILD && warp >= 2) {
    // ===== builders: d(conv2 out) of image i straight into the MN-major B operand stage =====
    for (int i = 0; i < n_img; ++i) {
      const int s = i & 1;
      if (i > 0) fetch_windows(blockIdx.x + i * gridDim.x);
      mbar_wait(&empty[s], ((i >> 1) & 1) ^ 1);
      T* __restrict__ pl = reinterpret_cast<T*>(dy_s + s * WG_DY_STAGE);
#pragma unroll
      for (int j = 0; j < NWIN; ++j) {
        const int wi = bt + j * NBUILD;
        if (wi < 32 * 49) {
          const int c = wi / 49, r = wi % 49, ph = r / 7, pw = r % 7;
          const float mu = s_bn2[c][0], is = s_bn2[c][1], g = s_bn2[c][2], b = s_bn2[c][3];
          bool live;
          const int arg = pool_winner<T>(yw[j], mu, is, g, b, live);
          const float d = live ? wd[j] : 0.f;
          const float k1 = s_k1[c], k2 = s_k2[c], k3 = is * g;
          const int po = (((c >> 3) * PL_PIX) + (2 * ph + 3) * 16 + 2 * pw) * 8 + (c & 7);
#pragma unroll
          for (int q4 = 0; q4 < 4; ++q4) {
            const float da = (live && q4 == arg) ? d : 0.f;
            stf<T>(pl + po + (q4 >> 1) * 128 + (q4 & 1) * 8, rnd<T>((da - k1 - (yw[j][q4] - mu) * k2) * k3));
          }
        }
      }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes -> visible to UMMA
      __syncwarp();
      if (lane == 0) mbar_arrive(&full[s]);
    }
  }
  if (warp == 4 || warp == 5) {  // TMEM lane quarters 0 and 1
    // epilogue: accumulator row m = kw*8 + ci%8 (m < 40: TMEM lanes 0..39), column = co
    mbar_wait(done, 0);
    tc_fence_after();
    const int q = warp & 3, m = q * 32 + lane, kw = m >> 3;
    float* row = a.gpart + (size_t)blockIdx.x * N_PAD;
#pragma unroll 1
    for (int acc = 0; acc < 2 * kh_n; ++acc) {
      const int kh = kh_lo + (acc >> 1), ci = (acc & 1) * 8 + (m & 7);
      float v[32];
      tmem_ld_row<32>(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * 32), v);
      if (kw < 5) {
#pragma unroll
        for (int co = 0; co < 32; ++co) row[OFF_W2 + (co * 16 + ci) * 25 + kh * 5 + kw] = v[co];
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(tmem_cols) : "memory");
    if (threadIdx.x == 0) TRACE(8 * 16 + 15);
  }
}

}  // namespace dstep
