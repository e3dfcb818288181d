// C ABI of libddpstep.so: context, launch sequence, CUDA-graph replay, peer bootstrap.
// Declarations and the reference call each entry point stands in for: include/ddpstep.h
#include "../../include/ddpstep.h"

#include <cstdarg>
#include <cstddef>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <type_traits>
#include <mutex>
#include <string>
#include <tuple>
#include <vector>

#include "bwd_kernels.cuh"
#include "conv2_tc.cuh"
#include "conv1_tc.cuh"
#include "bwd_mid_tc.cuh"
#include "fwd_kernels.cuh"
#include "step_kernel.cuh"

using namespace dstep;

static thread_local std::string g_create_error;

// mode: 0 train, 1 eval, 2 train micro-step inside no_sync(); source: 0 float batch on the device, 1 uint8 dataset + indices, 2 pinned host batch
typedef std::tuple<int, int, int, int> GraphKey;  // (mode, source, has_labels, batch)

struct GraphEntry {
  cudaGraphExec_t exec = nullptr;
  cudaGraph_t graph = nullptr;
  // the kernel nodes whose argument struct holds the input / label pointers, and host
  // copies of those structs: a new (x, y) pair only re-points these nodes of the
  // instantiated graph (cudaGraphExecKernelNodeSetParams), it does not re-capture
  cudaGraphNode_t node_c1 = nullptr, node_fc = nullptr, node_c1w = nullptr;
  Conv1Args a_c1;
  FcArgs a_fc;
  Conv1WgradArgs a_c1w;
  Conv1TcArgs a_c1tc;
  bool c1tc = false;  // the first input-consuming node is the tensor-core conv1 (large batch)
  Conv1WgradTcArgs a_c1wtc;
  bool c1wtc = false;
  BatchSrc x = {nullptr, nullptr, nullptr};
  const int64_t* y = nullptr;
  int n_kernels = 0;
  bool planes = false;  // p1 / d(conv2 out) exist only as padded planes after this graph ran (ddpstep_read_activation)
  // host-fed variant (ddpstep_train_step_host): the two H2D copies are nodes of the graph
  cudaGraphNode_t node_hx = nullptr, node_hy = nullptr;
  const void *hx = nullptr, *hy = nullptr;
};

struct ddpstep_ctx {
  ddpstep_config cfg;
  std::string err;
  // state
  float *params = nullptr, *grads = nullptr, *grads_scaled = nullptr, *bnbuf = nullptr;
  long long* nbt = nullptr;
  DevState* ds = nullptr;
  BnStat* st = nullptr;  // [2]
  float* counts = nullptr;
  // activations (sized by batch_cap, element size by dtype)
  void *y1 = nullptr, *p1 = nullptr, *y2 = nullptr, *p2 = nullptr, *logits = nullptr, *dlogits = nullptr;
  void *dflat = nullptr, *dy2 = nullptr, *dp1 = nullptr;
  void *p1h = nullptr, *dy2h = nullptr, *w2_tiles = nullptr;  // padded-plane copies for the tcgen05 conv2 path (zero halos, written once at alloc)
  bool use_tc = false, use_wgrad_tc = false, use_pdl = true, last_planes = false, use_conv1_tc = true, use_conv1_wgrad_tc = true;
  // small grids (<= 64 partial rows): 2 = tagged rows, the BN consumer's first CTA publishes the statistics before its grid
  // dependency resolves; 1 = plain rows, every consumer CTA reduces them; 0 = the producer's last CTA finalises (DDPSTEP_DEFER)
  int defer_stats = 2;
  // BN2 backward apply + conv2 data gradient + BN1 backward pass 1 as one kernel (bwd_mid_tc.cuh).  Opt-in
  // (DDPSTEP_MID=1): measured 47.0 vs 45.2 us per step -- the fused kernel itself saves ~4 us on the main chain, but the
  // conv2 weight-gradient branch can then only start after it and becomes the critical path.
  bool fuse_mid = false;
  bool fuse_head = true;  // BN2 + ReLU + pool inside the Linear+CE kernel at small batch (DDPSTEP_HEAD=0: separate kernel)
  float *nll = nullptr, *part1 = nullptr, *part2 = nullptr, *bpart = nullptr, *gpart = nullptr, *gdirect = nullptr;
  // the same partial rows as tagged 8-byte words (store_partial), one buffer per producer so that a tag can never match
  // another stage's words of the same step
  float *part1_ll = nullptr, *part2_ll = nullptr, *bpart1_ll = nullptr, *bpart2_ll = nullptr;
  float* x_stage = nullptr;
  int64_t* y_stage = nullptr;
  // comm
  SymRegion* sym = nullptr;
  CommDev comm;
  bool peers_ready = false;
  std::vector<void*> opened;
  // one captured graph per (mode, where the batch comes from, labels present, batch size)
  std::map<GraphKey, GraphEntry> graphs;
  // (source kind, n_steps, batch, dataset pointer, label pointer) -> K training steps captured as ONE graph
  std::map<std::tuple<int, int, int, const void*, const void*>, GraphEntry> multi_graphs;
  cudaStream_t cap_stream = nullptr, side1 = nullptr, side2 = nullptr;
  cudaEvent_t ev_f0 = nullptr, ev_j0 = nullptr, ev_f1 = nullptr, ev_j1 = nullptr, ev_f2 = nullptr, ev_j2 = nullptr;
  cudaEvent_t ev_h0 = nullptr, ev_h1 = nullptr, ev_m = nullptr;
  float* gred = nullptr;
  float* mom = nullptr;   // SGD momentum buffers, same flat layout as the parameters
  llword* tail_flags = nullptr;  // per-CTA flag words of the grid-shaped tail kernel
  bool tail_grid = true;  // tail kernel as 29 ordinary CTAs (default) or as one 8-CTA cluster (DDPSTEP_TAIL=cluster)
  float* acc = nullptr;   // gradient accumulated by DDP.no_sync() micro-steps, same layout
  bool accumulate = false;  // the step being enqueued is such a micro-step
  volatile float* host_mail = nullptr;  // pinned, mapped: {loss, completed-step counter}
  unsigned int steps_enqueued = 0;     // training steps launched so far (matches DevState.epoch once they finish)
  bool early_push = true;              // world > 1: side branches push their gradient slices (DDPSTEP_EARLY_PUSH=0: the tail does)
  bool chained_step = false;           // the step being enqueued follows another one inside the same captured graph
  float* xs_stage = nullptr; int64_t* ys_stage = nullptr; long long* idx_stage = nullptr;  // multi-step staging (lazy)
  size_t multi_cap = 0;                // images the multi-step staging buffers hold
  DevState::SgdGroup groups[10];       // host copies of the per-tensor SGD settings (source of the async uploads)
  bool post_loss = false;              // the step being enqueued posts its loss to the pinned mailbox (host entry point)
  int sm_khz = 1965000;                // SM clock used to turn the peer timeout into clock64() cycles
  StepArgs last_step;                  // arguments of the most recently enqueued tail kernel (ddpstep_allreduce_only)
  PushArgs last_push_fc, last_push_c2; // ... and of the two early-push kernels (world > 1)
  int last_c2_ctas = 25;
  int last_step_dtype_ok = 0;
  int cap_bn1 = 0, cap_bn2 = 0, cap_bwd2 = 0, cap_c1w = 0;  // CTAs of the BN consumer kernels the device holds at once (occupancy API)
  int64_t launches = 0;
  int esize = 2;
};

static int fail(ddpstep_ctx* c, int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  if (c) c->err = buf; else g_create_error = buf;
  return code;
}

#define CK(call)                                                                                   \
  do {                                                                                             \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess)                                                                         \
      return fail(ctx, DDPSTEP_E_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
  } while (0)

template <typename P> static int dalloc(ddpstep_ctx* ctx, P** p, size_t bytes) {
  CK(cudaMalloc(reinterpret_cast<void**>(p), bytes));
  CK(cudaMemset(*p, 0, bytes));
  return 0;
}

extern "C" void ddpstep_default_config(ddpstep_config* c) {
  memset(c, 0, sizeof *c);
  c->abi_version = DDPSTEP_ABI_VERSION;
  c->rank = 0; c->world = 1; c->device = 0;
  c->dtype = DDPSTEP_FP16;
  c->batch_cap = 32;
  c->use_scaler = 1; c->sync_bn = 1; c->use_graph = 1; c->conv_bias_double_round = 1;
  c->lr = 1e-4f;
  c->init_scale = 65536.0f; c->growth_factor = 2.0f; c->backoff_factor = 0.5f; c->growth_interval = 2000;
  c->bn_momentum = 0.1f; c->bn_eps = 1e-5f;
  c->momentum = 0.f; c->weight_decay = 0.f; c->dampening = 0.f; c->nesterov = 0;
}

extern "C" const char* ddpstep_last_error(const ddpstep_ctx* ctx) {
  return ctx ? ctx->err.c_str() : g_create_error.c_str();
}

extern "C" int ddpstep_create(const ddpstep_config* cfg, ddpstep_ctx** out) {
  ddpstep_ctx* ctx = nullptr;
  if (!cfg || !out) return fail(nullptr, DDPSTEP_E_ARG, "null argument");
  if (cfg->abi_version != DDPSTEP_ABI_VERSION) return fail(nullptr, DDPSTEP_E_ARG, "ABI version mismatch");
  if (cfg->world < 1 || cfg->world > MAX_WORLD || cfg->rank < 0 || cfg->rank >= cfg->world)
    return fail(nullptr, DDPSTEP_E_ARG, "bad rank/world %d/%d (world <= %d)", cfg->rank, cfg->world, MAX_WORLD);
  if (cfg->batch_cap < 1) return fail(nullptr, DDPSTEP_E_ARG, "batch_cap must be >= 1");
  if (cfg->dtype < 0 || cfg->dtype > 2) return fail(nullptr, DDPSTEP_E_ARG, "bad dtype");
  if (cfg->momentum < 0.f || cfg->weight_decay < 0.f) return fail(nullptr, DDPSTEP_E_ARG, "momentum and weight_decay must be >= 0");
  if (cfg->nesterov && (cfg->momentum <= 0.f || cfg->dampening != 0.f))
    return fail(nullptr, DDPSTEP_E_ARG, "nesterov momentum requires a momentum and zero dampening");  // sgd.py:68
  CK(cudaSetDevice(cfg->device));
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, cfg->device));
  if (prop.major != 10)
    return fail(nullptr, DDPSTEP_E_ARCH, "device %d is sm_%d%d; this library is built for sm_100a only", cfg->device,
                prop.major, prop.minor);
  ctx = new ddpstep_ctx();
  ctx->cfg = *cfg;
  ctx->esize = cfg->dtype == DDPSTEP_FP32 ? 4 : 2;
  const size_t B = (size_t)cfg->batch_cap, es = (size_t)ctx->esize;
  int rc = 0;
#define A(p, bytes) if (!rc) rc = dalloc(ctx, &ctx->p, (bytes))
  A(params, N_PAD * 4); A(grads, N_PAD * 4); A(grads_scaled, N_PAD * 4); A(bnbuf, N_BNBUF * 4);
  A(nbt, 2 * 8); A(ds, sizeof(DevState)); A(st, 2 * sizeof(BnStat)); A(counts, 2 * 4);
  A(y1, B * 16 * 784 * es); A(p1, B * 16 * 196 * es); A(y2, B * 32 * 196 * es); A(p2, B * 1568 * es);
  A(logits, B * 10 * es); A(dlogits, B * 10 * es); A(dflat, B * 1568 * es); A(dy2, B * 32 * 196 * es);
  A(dp1, B * 16 * 196 * es); A(nll, B * 4);
  A(part1, 296 * 16 * 2 * 4); A(part2, 296 * 32 * 2 * 4); A(bpart, 32 * 32 * 2 * 4);
  A(part1_ll, 64 * 32 * 8); A(part2_ll, 64 * 64 * 8); A(bpart1_ll, 16 * 64 * 2 * 8); A(bpart2_ll, 32 * 64 * 2 * 8);
  A(gpart, (size_t)MAX_PART_ROWS * N_PAD * 4); A(gdirect, N_PAD * 4);
  A(x_stage, B * 784 * 4); A(y_stage, B * 8); A(gred, N_PAD * 4); A(mom, N_PAD * 4); A(acc, N_PAD * 4); A(tail_flags, 64 * 8);
  {
    const char* env = getenv("DDPSTEP_TC");
    ctx->use_tc = cfg->dtype != DDPSTEP_FP32 && !(env && env[0] == '0');
    const char* envcw = getenv("DDPSTEP_CONV1_WGRAD_TC");
    ctx->use_conv1_wgrad_tc = !(envcw && envcw[0] == '0');
    const char* envc = getenv("DDPSTEP_CONV1_TC");
    ctx->use_conv1_tc = !(envc && envc[0] == '0');
    const char* envp = getenv("DDPSTEP_PDL");
    ctx->use_pdl = !(envp && envp[0] == '0');
    const char* envd = getenv("DDPSTEP_DEFER");
    if (envd && envd[0] >= '0' && envd[0] <= '2') ctx->defer_stats = envd[0] - '0';
    const char* envt = getenv("DDPSTEP_TAIL");
    ctx->tail_grid = !(envt && envt[0] == 'c');
    const char* enve = getenv("DDPSTEP_EARLY_PUSH");
    ctx->early_push = !(enve && enve[0] == '0');
    const char* envm = getenv("DDPSTEP_MID");
    ctx->fuse_mid = envm && envm[0] == '1';
    const char* envh = getenv("DDPSTEP_HEAD");
    ctx->fuse_head = !(envh && envh[0] == '0');
    const char* envw = getenv("DDPSTEP_WGRAD_TC");
    ctx->use_wgrad_tc = ctx->use_tc && !(envw && envw[0] == '0');
  }
  if (ctx->use_tc) { A(p1h, B * 2 * PL_BYTES); A(dy2h, B * 4 * PL_BYTES); A(w2_tiles, 2 * TC_WBYTES); }
  A(sym, sizeof(SymRegion));
#undef A
  if (rc) { g_create_error = ctx->err; ddpstep_destroy(ctx); return rc; }
  // running_var starts at 1 (nn.BatchNorm2d), scaler at init_scale (grad_scaler.py:173-179)
  std::vector<float> bn(N_BNBUF, 0.f);
  for (int i = 0; i < 16; ++i) bn[BNB_RV1 + i] = 1.f;
  for (int i = 0; i < 32; ++i) bn[BNB_RV2 + i] = 1.f;
  DevState h;
  memset(&h, 0, sizeof h);
  h.scale = cfg->use_scaler ? cfg->init_scale : 1.0f;
  for (int t = 0; t < 10; ++t) h.grp[t] = {cfg->lr, cfg->momentum, cfg->weight_decay, cfg->dampening, cfg->nesterov};
  ctx->sm_khz = prop.clockRate > 0 ? prop.clockRate : 1965000;
  {
    // default peer timeout: 20 s (a rank may legitimately be seconds late: first-batch DataLoader start, a rank-0
    // checkpoint, a graph capture); DDPSTEP_PEER_TIMEOUT_MS or ddpstep_set_peer_timeout_ms change it
    const char* et = getenv("DDPSTEP_PEER_TIMEOUT_MS");
    const double ms = et ? atof(et) : 20000.0;
    h.timeout_cycles = (long long)(ms * (double)ctx->sm_khz);
  }
  cudaError_t e = cudaMemcpy(ctx->bnbuf, bn.data(), N_BNBUF * 4, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMemcpy(ctx->ds, &h, sizeof h, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&ctx->cap_stream, cudaStreamNonBlocking);
  if (e == cudaSuccess) { void* hm = nullptr; e = cudaHostAlloc(&hm, 64, cudaHostAllocMapped); if (e == cudaSuccess) { memset(hm, 0, 64); ctx->host_mail = static_cast<volatile float*>(hm); } }
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&ctx->side1, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&ctx->side2, cudaStreamNonBlocking);
  for (cudaEvent_t* ev : {&ctx->ev_f0, &ctx->ev_j0, &ctx->ev_f1, &ctx->ev_j1, &ctx->ev_f2, &ctx->ev_j2, &ctx->ev_h0, &ctx->ev_h1, &ctx->ev_m})
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(ev, cudaEventDisableTiming);
  // kernels that need more than 48 KB of dynamic shared memory
  const int smem_dgrad = (32 * 18 * C14_RS + 8 * 32 * 25) * 4;
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k_conv14<float, 32, 16, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_dgrad);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k_conv14<__half, 32, 16, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_dgrad);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k_conv14<__nv_bfloat16, 32, 16, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_dgrad);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k_conv2_wgrad_tc<__half, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, WG_SMEM_BYTES);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k_conv2_wgrad_tc<__nv_bfloat16, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, WG_SMEM_BYTES);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k_conv2_wgrad_tc<__half, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, WG_SMEM_BYTES);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k_conv2_wgrad_tc<__nv_bfloat16, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, WG_SMEM_BYTES);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k_conv1_wgrad_tc<__half>, cudaFuncAttributeMaxDynamicSharedMemorySize, C1W_TC_SMEM);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k_conv1_wgrad_tc<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, C1W_TC_SMEM);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k_bwd_mid_tc<__half>, cudaFuncAttributeMaxDynamicSharedMemorySize, MID_SMEM_BYTES);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k_bwd_mid_tc<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, MID_SMEM_BYTES);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k_conv1_tc<__half>, cudaFuncAttributeMaxDynamicSharedMemorySize, C1T_SMEM);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k_conv1_tc<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, C1T_SMEM);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k_conv2_tc<__half, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, tc_smem_bytes(2));
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k_conv2_tc<__nv_bfloat16, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, tc_smem_bytes(2));
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k_conv2_tc<__half, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, tc_smem_bytes(4));
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k_conv2_tc<__nv_bfloat16, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, tc_smem_bytes(4));
  if (e == cudaSuccess) {
    // Early publication (DDPSTEP_DEFER=2): the CTAs of a BN consumer grid spin on words that CTA 0 of the
    // SAME grid publishes.  CUDA promises no dispatch order, so the scheme is only used for grids that are resident
    // all at once according to the occupancy API (checked per launch in enqueue_train); larger grids let every
    // consumer CTA reduce the rows itself (the DEFER=1 scheme).
    int occ[3] = {0, 0, 0}, occ_c1w = 0;
    const void* f[3];
    switch (cfg->dtype) {
      case DDPSTEP_FP32: f[0] = (const void*)k_bn_relu_pool<float, 16, 28>; f[1] = (const void*)k_bn_relu_pool<float, 32, 14>; f[2] = (const void*)k_bn2_bwd_apply<float>; break;
      case DDPSTEP_FP16: f[0] = (const void*)k_bn_relu_pool<__half, 16, 28>; f[1] = (const void*)k_bn_relu_pool<__half, 32, 14>; f[2] = (const void*)k_bn2_bwd_apply<__half>; break;
      default: f[0] = (const void*)k_bn_relu_pool<__nv_bfloat16, 16, 28>; f[1] = (const void*)k_bn_relu_pool<__nv_bfloat16, 32, 14>; f[2] = (const void*)k_bn2_bwd_apply<__nv_bfloat16>; break;
    }
    for (int i = 0; i < 3; ++i)
      if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ[i], f[i], 256, 0);
    ctx->cap_bn1 = occ[0] * prop.multiProcessorCount;
    ctx->cap_bn2 = occ[1] * prop.multiProcessorCount;
    ctx->cap_bwd2 = occ[2] * prop.multiProcessorCount;
    const void* fw = cfg->dtype == DDPSTEP_FP32 ? (const void*)k_conv1_wgrad<float>
                   : cfg->dtype == DDPSTEP_FP16 ? (const void*)k_conv1_wgrad<__half> : (const void*)k_conv1_wgrad<__nv_bfloat16>;
    if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_c1w, fw, C1W_THREADS, 0);
    ctx->cap_c1w = occ_c1w * prop.multiProcessorCount;
  }
  if (e != cudaSuccess) {
    fail(nullptr, DDPSTEP_E_CUDA, "context setup failed: %s", cudaGetErrorString(e));
    ddpstep_destroy(ctx);
    return DDPSTEP_E_CUDA;
  }
  memset(&ctx->comm, 0, sizeof ctx->comm);
  ctx->comm.rank = cfg->rank;
  ctx->comm.world = cfg->world;
  ctx->comm.self = ctx->sym;
  ctx->comm.peer[cfg->rank] = ctx->sym;
  ctx->peers_ready = cfg->world == 1;
  *out = ctx;
  return 0;
}

static void drop_graphs(ddpstep_ctx* ctx) {
  for (auto& kv : ctx->graphs) {
    if (kv.second.exec) cudaGraphExecDestroy(kv.second.exec);
    if (kv.second.graph) cudaGraphDestroy(kv.second.graph);
  }
  ctx->graphs.clear();
  for (auto& kv : ctx->multi_graphs) {
    if (kv.second.exec) cudaGraphExecDestroy(kv.second.exec);
    if (kv.second.graph) cudaGraphDestroy(kv.second.graph);
  }
  ctx->multi_graphs.clear();
}

extern "C" int ddpstep_destroy(ddpstep_ctx* ctx) {
  if (!ctx) return 0;
  cudaSetDevice(ctx->cfg.device);
  cudaDeviceSynchronize();
  drop_graphs(ctx);
  for (void* p : ctx->opened) cudaIpcCloseMemHandle(p);
  void* bufs[] = {ctx->params, ctx->grads, ctx->grads_scaled, ctx->bnbuf, ctx->nbt, ctx->ds, ctx->st, ctx->counts,
                  ctx->y1, ctx->p1, ctx->y2, ctx->p2, ctx->logits, ctx->dlogits, ctx->dflat, ctx->dy2, ctx->dp1,
                  ctx->nll, ctx->part1, ctx->part2, ctx->bpart, ctx->gpart, ctx->gdirect, ctx->x_stage,
                  ctx->y_stage, ctx->sym, ctx->p1h, ctx->dy2h, ctx->w2_tiles, ctx->gred,
                  ctx->part1_ll, ctx->part2_ll, ctx->bpart1_ll, ctx->bpart2_ll, ctx->mom, ctx->acc,
                  ctx->xs_stage, ctx->ys_stage, ctx->idx_stage, ctx->tail_flags};
  for (void* p : bufs) if (p) cudaFree(p);
  if (ctx->host_mail) cudaFreeHost(const_cast<float*>(ctx->host_mail));
  if (ctx->cap_stream) cudaStreamDestroy(ctx->cap_stream);
  if (ctx->side1) cudaStreamDestroy(ctx->side1);
  if (ctx->side2) cudaStreamDestroy(ctx->side2);
  for (cudaEvent_t ev : {ctx->ev_f0, ctx->ev_j0, ctx->ev_f1, ctx->ev_j1, ctx->ev_f2, ctx->ev_j2, ctx->ev_h0, ctx->ev_h1, ctx->ev_m}) if (ev) cudaEventDestroy(ev);
  delete ctx;
  return 0;
}

extern "C" int ddpstep_export_handle(ddpstep_ctx* ctx, void* handle64) {
  if (!ctx || !handle64) return fail(ctx, DDPSTEP_E_ARG, "null argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == DDPSTEP_HANDLE_BYTES, "handle size");
  CK(cudaSetDevice(ctx->cfg.device));
  cudaIpcMemHandle_t h;
  CK(cudaIpcGetMemHandle(&h, ctx->sym));
  memcpy(handle64, &h, sizeof h);
  return 0;
}

extern "C" int ddpstep_import_peers(ddpstep_ctx* ctx, const void* handles, int32_t world) {
  if (!ctx || !handles) return fail(ctx, DDPSTEP_E_ARG, "null argument");
  if (world != ctx->cfg.world) return fail(ctx, DDPSTEP_E_ARG, "world %d != configured %d", world, ctx->cfg.world);
  if (ctx->peers_ready) return fail(ctx, DDPSTEP_E_STATE, "peers already imported");
  CK(cudaSetDevice(ctx->cfg.device));
  for (int r = 0; r < world; ++r) {
    if (r == ctx->cfg.rank) continue;
    cudaIpcMemHandle_t h;
    memcpy(&h, static_cast<const char*>(handles) + (size_t)r * DDPSTEP_HANDLE_BYTES, sizeof h);
    void* p = nullptr;
    CK(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    ctx->opened.push_back(p);
    ctx->comm.peer[r] = static_cast<SymRegion*>(p);
  }
  ctx->peers_ready = true;
  return 0;
}

extern "C" int ddpstep_load_state(ddpstep_ctx* ctx, const float* params, const float* bn, const int64_t* nbt) {
  if (!ctx || !params || !bn) return fail(ctx, DDPSTEP_E_ARG, "null argument");
  CK(cudaSetDevice(ctx->cfg.device));
  CK(cudaDeviceSynchronize());
  CK(cudaMemcpy(ctx->params, params, N_PARAM * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(ctx->bnbuf, bn, N_BNBUF * 4, cudaMemcpyHostToDevice));
  int64_t z[2] = {0, 0};
  CK(cudaMemcpy(ctx->nbt, nbt ? nbt : z, 16, cudaMemcpyHostToDevice));
  if (ctx->use_tc) {
    if (ctx->cfg.dtype == DDPSTEP_FP16) k_refresh_w2_tiles<__half><<<100, 256>>>(ctx->params, ctx->w2_tiles);
    else k_refresh_w2_tiles<__nv_bfloat16><<<100, 256>>>(ctx->params, ctx->w2_tiles);
    CK(cudaGetLastError());
    CK(cudaDeviceSynchronize());
  }
  return 0;
}

extern "C" int ddpstep_read_state(ddpstep_ctx* ctx, float* params, float* bn, int64_t* nbt, float* scale,
                                  int32_t* tracker) {
  if (!ctx) return fail(ctx, DDPSTEP_E_ARG, "null argument");
  CK(cudaSetDevice(ctx->cfg.device));
  CK(cudaDeviceSynchronize());
  if (params) CK(cudaMemcpy(params, ctx->params, N_PARAM * 4, cudaMemcpyDeviceToHost));
  if (bn) CK(cudaMemcpy(bn, ctx->bnbuf, N_BNBUF * 4, cudaMemcpyDeviceToHost));
  if (nbt) CK(cudaMemcpy(nbt, ctx->nbt, 16, cudaMemcpyDeviceToHost));
  DevState h;
  CK(cudaMemcpy(&h, ctx->ds, sizeof h, cudaMemcpyDeviceToHost));
  if (scale) *scale = h.scale;
  if (tracker) *tracker = h.growth_tracker;
  return 0;
}

extern "C" int ddpstep_set_scaler(ddpstep_ctx* ctx, float scale, int32_t tracker) {
  if (!ctx) return fail(ctx, DDPSTEP_E_ARG, "null argument");
  CK(cudaSetDevice(ctx->cfg.device));
  CK(cudaDeviceSynchronize());
  CK(cudaMemcpy(&ctx->ds->scale, &scale, 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(&ctx->ds->growth_tracker, &tracker, 4, cudaMemcpyHostToDevice));
  return 0;
}

extern "C" int ddpstep_state_ptrs(ddpstep_ctx* ctx, ddpstep_ptrs* o) {
  if (!ctx || !o) return fail(ctx, DDPSTEP_E_ARG, "null argument");
  o->params = ctx->params; o->grads = ctx->grads; o->grads_scaled = ctx->grads_scaled;
  o->bn_buffers = ctx->bnbuf; o->num_batches_tracked = reinterpret_cast<int64_t*>(ctx->nbt);
  o->scale = &ctx->ds->scale; o->growth_tracker = &ctx->ds->growth_tracker;
  o->found_inf = &ctx->ds->found_inf; o->loss = &ctx->ds->loss;
  o->logits = ctx->logits; o->counts = ctx->counts;
  o->input_stage = ctx->x_stage; o->label_stage = ctx->y_stage;
  o->momentum_buffer = ctx->mom; o->momentum_ready = &ctx->ds->mom_ready;
  return 0;
}

// ------------------------------------------------------------------------------------ launch sequences

static inline int cdiv(int a, int b) { return (a + b - 1) / b; }
static inline int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

struct Recorder {  // keeps the argument structs that hold x / y so graph replays can be re-pointed
  int n = 0;
  std::vector<cudaEvent_t>* events = nullptr;  // profiling: one event after every kernel
  cudaStream_t ev_stream = nullptr;
  bool ev_external = false;
  unsigned int skipped = 0;  // bit i: profile slot i belongs to a kernel that did not run (its work rides in another one)
  Conv1Args a_c1;
  FcArgs a_fc;
  Conv1WgradArgs a_c1w;
  Conv1TcArgs a_c1tc;
  bool c1tc = false;
  Conv1WgradTcArgs a_c1wtc;
  bool c1wtc = false;
};

// Kernel launch with an optional programmatic (PDL) edge to the previous kernel of the stream:
// the kernel may start while its predecessor is still running and orders itself with pdl_wait().
template <typename A>
static void launch_k(void (*kernel)(A), dim3 grid, dim3 block, size_t smem, cudaStream_t s, const A& args, bool pdl) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
  cudaLaunchKernelEx(&cfg, kernel, args);
}

template <typename T>
static cudaError_t enqueue_train(ddpstep_ctx* c, const BatchSrc& x, const int64_t* y, int B, cudaStream_t s, Recorder* rec) {
  const ddpstep_config& cfg = c->cfg;
  const int dbl = cfg.conv_bias_double_round, sync = cfg.sync_bn;
  BnStat* st1 = c->st;
  BnStat* st2 = c->st + 1;
  auto mark = [&]() {
    if (rec) {
      rec->n++;
      if (rec->events) {
        cudaEvent_t ev;
        cudaEventCreate(&ev);
        if (rec->ev_external) cudaEventRecordWithFlags(ev, rec->ev_stream, cudaEventRecordExternal);
        else cudaEventRecord(ev, rec->ev_stream);
        rec->events->push_back(ev);
      }
    }
    c->launches++;
  };
  auto skip_slot = [&]() {  // keeps the profile's 15 slots aligned: an empty slot for a kernel that did not run
    if (rec && rec->events) {
      rec->skipped |= 1u << rec->events->size();
      cudaEvent_t ev;
      cudaEventCreate(&ev);
      if (rec->ev_external) cudaEventRecordWithFlags(ev, rec->ev_stream, cudaEventRecordExternal);
      else cudaEventRecord(ev, rec->ev_stream);
      rec->events->push_back(ev);
    }
  };
  // the two weight-gradient branches and the filter-tile refresh do not sit on the activation chain:
  // they run on side streams (parallel branches of the captured graph)
  const bool fork = !(rec && rec->events);  // per-kernel profiling keeps everything on one stream
  const bool pdl = fork && c->use_pdl;
  cudaStream_t s1 = fork ? c->side1 : s, s2 = fork ? c->side2 : s;
  // 0. conv2 filter -> tcgen05 B tiles.  Small batch: extra CTAs of the conv1 kernel build them (no extra node, conv2
  //    keeps its programmatic edge).  Large batch (conv1 on the tensor cores): a side branch that joins before conv2.
  bool tiles = false, tiles_in_conv1 = false;
  if constexpr (!std::is_same<T, float>::value) {
    tiles = c->use_tc;
    tiles_in_conv1 = tiles && !(c->use_conv1_tc && B > 64);
  }
  if (tiles && !tiles_in_conv1) {
    if (fork) { cudaEventRecord(c->ev_f0, s); cudaStreamWaitEvent(s1, c->ev_f0, 0); }
    if constexpr (!std::is_same<T, float>::value) k_refresh_w2_tiles<T><<<100, 256, 0, s1>>>(c->params, c->w2_tiles);
    mark();
    if (fork) cudaEventRecord(c->ev_j0, s1);
  } else {
    skip_slot();
  }
  // 1. conv1 + local BN1 statistics
  const int g1 = std::min(B, 296);
  // small grids: the BN kernels reduce the conv kernels' partial sums themselves (no last-CTA tail)
  const int tsplit = B <= 64 ? 2 : 1;  // tcgen05 conv2: one M-tile per CTA at small batch
  // conv2 on the tensor cores writes one partial row per (image CTA, M-tile): 2B rows at small batch.  The tagged-row
  // scheme reads at most 64 rows; between 33 and 64 images the plain deferred scheme (any row count) takes over.
  const int rows2 = std::min(B, 296) * ((!std::is_same<T, float>::value && c->use_tc) ? tsplit : 1);
  // tagged rows + early publication only for consumer grids that are fully co-resident (see ddpstep_create)
  const int grid_bn1 = clampi(cdiv(B * 16 * 196, 256), 1, 1184), grid_bn2 = clampi(cdiv(B * 32 * 49, 256), 1, 1184);
  int defer1 = g1 <= 64 ? c->defer_stats : 0;
  if (defer1 == 2 && grid_bn1 > c->cap_bn1) defer1 = 1;
  int defer2 = std::min(B, 296) > 64 ? 0 : (rows2 <= 64 ? c->defer_stats : std::min(c->defer_stats, 1));
  if (defer2 == 2 && grid_bn2 > c->cap_bn2) defer2 = 1;
  // the same rule for the two backward reductions: their consumers are k_bn2_bwd_apply(_planes) (<= 512 CTAs) and the
  // conv1 weight-gradient kernel (<= 1184 CTAs of 160 threads on the CUDA-core path, <= 296 on the tensor-core one)
  const int dstat2 = (c->defer_stats == 2 && 32 * 16 > c->cap_bwd2) ? 1 : c->defer_stats;
  const bool cuda_core_c1w = std::is_same<T, float>::value || !(c->use_tc && c->use_conv1_tc && c->use_conv1_wgrad_tc && B > 64);
  const int dstat1 = (c->defer_stats == 2 && cuda_core_c1w && std::min(B, 296) * 4 > c->cap_c1w) ? 1 : c->defer_stats;
  Conv1Args a1{x, c->params, c->y1, defer1 == 2 ? c->part1_ll : c->part1, defer1, st1, c->ds, c->comm, B, dbl, sync, cfg.bn_eps,
               tiles_in_conv1 ? c->w2_tiles : nullptr, g1};
  constexpr int kTileCtas = 25600 / (C1_GROUPS * C1_THREADS);  // 50 extra blockIdx.x values x 4 groups x 128 threads
  if (rec) rec->a_c1 = a1;
  bool c1tc = false;
  int rows1 = g1;
  if constexpr (!std::is_same<T, float>::value) {
    if (c->use_tc && c->use_conv1_tc && B > 64) {  // large batch: conv1 on the tensor cores (conv1_tc.cuh)
      c1tc = true;
      rows1 = std::min(B, 296);  // two CTAs per SM
      Conv1TcArgs t1{x, c->params, c->y1, c->part1, 0, st1, c->ds, c->comm, B, dbl, sync, cfg.bn_eps};
      if (rec) { rec->a_c1tc = t1; rec->c1tc = true; }
      k_conv1_tc<T><<<rows1, C1T_THREADS, C1T_SMEM, s>>>(t1);
    }
  }
  // in a multi-step graph the conv1 of step s > 0 hangs off the previous step's tail kernel by a programmatic edge
  if (!c1tc) launch_k(k_conv1_fwd<T>, dim3(g1 + (tiles_in_conv1 ? kTileCtas : 0), C1_GROUPS), dim3(C1_THREADS), 0, s, a1, pdl && c->chained_step);
  mark();
  // 2. BN1 + ReLU + pool
  BnPoolArgs b1{c->y1, c->p1, c->use_tc ? c->p1h : nullptr, c->params, st1, c->bnbuf, c->nbt, c->ds, c->comm, B, 0, 0, sync, cfg.bn_momentum, cfg.bn_eps,
                defer1 == 2 ? c->part1_ll : defer1 ? c->part1 : nullptr, rows1, defer1 == 2};
  bool planes = false;
  // large batches: write p1 / d(conv2 out) only as padded planes with 16-byte stores (nobody reads the NCHW copies);
  // small batches keep the one-thread-per-element kernels, whose shorter per-thread chains give lower latency
  if constexpr (!std::is_same<T, float>::value) planes = c->use_tc && c->use_wgrad_tc && B > 64;
  c->last_planes = planes;
  if (planes) {
    if constexpr (!std::is_same<T, float>::value)
      launch_k(k_bn1_relu_pool_planes<T>, dim3(clampi(cdiv(B * 2 * 196, 256), 1, 1184)), dim3(256), 0, s, b1, pdl);
  } else {
    launch_k(k_bn_relu_pool<T, 16, 28>, dim3(clampi(cdiv(B * 16 * 196, 256), 1, 1184)), dim3(256), 0, s, b1, pdl);
  }
  mark();
  // 3. conv2 + local BN2 statistics
  const int g2 = std::min(B, 296);  // tensor-core path: up to two CTAs per SM
  Conv14Args a2{c->p1, c->params, c->y2, defer2 == 2 ? c->part2_ll : c->part2, defer2, st2, c->ds, c->comm, B, dbl, sync, cfg.bn_eps};
  bool tc = false;
  if constexpr (!std::is_same<T, float>::value) {
    if (c->use_tc) {  // tcgen05 implicit GEMM on the padded planes written by kernel 2
      tc = true;
      if (fork && !tiles_in_conv1) cudaStreamWaitEvent(s, c->ev_j0, 0);
      Conv2TcArgs t2{c->w2_tiles, c->p1h, c->params, c->y2, defer2 == 2 ? c->part2_ll : c->part2, defer2, st2, c->ds, c->comm, B, dbl, sync, cfg.bn_eps};
      launch_k(k_conv2_tc<T, 0>, dim3(g2, tsplit), dim3(TC_THREADS), tc_smem_bytes(2), s, t2, pdl);
    }
  }
  if (!tc) k_conv14<T, 16, 32, 0><<<dim3(g2, 4), 128, (16 * 18 * C14_RS + 8 * 16 * 25) * 4, s>>>(a2);
  mark();
  // 4. BN2 + ReLU + pool
  BnPoolArgs b2{c->y2, c->p2, nullptr, c->params, st2, c->bnbuf, c->nbt, c->ds, c->comm, B, 1, 0, sync, cfg.bn_momentum, cfg.bn_eps,
                defer2 == 2 ? c->part2_ll : defer2 ? c->part2 : nullptr, tc ? g2 * tsplit : g2, defer2 == 2};
  // small batch: BN2's backward pass 1 rides in the Linear+CE kernel (one image per CTA, one tagged partial row per
  // image), and so does BN2 + ReLU + pool itself when conv2 left tagged rows (`head`: slot 4 stays empty)
  const bool fuse_bn2r = dstat2 == 2 && B <= 64;
  const bool head = fuse_bn2r && defer2 == 2 && c->fuse_head;
  if (!head) { launch_k(k_bn_relu_pool<T, 32, 14>, dim3(clampi(cdiv(B * 32 * 49, 256), 1, 1184)), dim3(256), 0, s, b2, pdl); mark(); }
  else skip_slot();
  // 5. Linear + CE + dlogits + d(pool2 out)
  FcArgs f{c->p2, c->params, reinterpret_cast<const long long*>(y), x.idx, c->logits, c->dlogits, c->dflat, c->nll, c->counts,
           c->ds, B, cfg.use_scaler, 0, c->y2, st2, fuse_bn2r ? c->bpart2_ll : nullptr, b2, head ? 1 : 0};
  if (rec) rec->a_fc = f;
  if (B > 1184) launch_k(k_fc_ce<T, 4>, dim3(296), dim3(224), 0, s, f, pdl);
  else launch_k(k_fc_ce<T, 1>, dim3(std::min(B, 296)), dim3(224), 0, s, f, pdl);
  mark();
  // 6. fc weight/bias gradient partials
  const int nsf = clampi(cdiv(B, 8), 1, 40);
  FcWgradArgs fw{c->p2, c->dlogits, c->gpart, B, cdiv(B, nsf)};
  if (fork) { cudaEventRecord(c->ev_f1, s); cudaStreamWaitEvent(s1, c->ev_f1, 0); }
  k_fc_wgrad<T><<<dim3(7, nsf), 224, 0, s1>>>(fw); mark();
  const bool early_on = cfg.world > 1 && c->early_push && !c->accumulate;  // a no_sync micro-step exchanges nothing
  if (early_on) {  // not a profile slot: it exists only at world > 1
    PushArgs pf{c->gpart, nullptr, nsf, 0, c->comm, c->ds, 1, c->acc};
    k_fc_reduce_push<T><<<cdiv(N_PAD - OFF_WFC, 512), 128, 0, s1>>>(pf);
    c->last_push_fc = pf;
    if (rec) rec->n++;
    c->launches++;
  }
  if (fork) cudaEventRecord(c->ev_j1, s1);
  // 7/8. BN2 backward: reduce, then d(conv2 out)
  const int ns2 = clampi(cdiv(B, 4), 1, 16);  // batch splits of the two BN2 backward kernels
  BnBwdArgs r2{c->y2, c->dflat, c->params, st2, dstat2 == 2 ? c->bpart2_ll : c->bpart, c->gdirect, c->ds, c->comm, B, cdiv(B, ns2), 1, sync, dstat2,
               fuse_bn2r ? B : ns2 /* partial rows per channel */, c->dy2, c->use_tc ? c->dy2h : nullptr, c->gpart};
  // the plane-writing apply kernel has 4 channel groups per batch split: use more splits to fill the GPU
  const int ns2p = clampi(cdiv(B, 2), 1, 64);
  BnBwdArgs r2p = r2;
  r2p.per_split = cdiv(B, ns2p);
  if (!fuse_bn2r) { launch_k(k_bn_bwd_reduce<T, 32, 14>, dim3(32, ns2), dim3(256), 0, s, r2, pdl); mark(); }
  else skip_slot();
  // small batch on the tensor-core path: BN2 backward apply, the conv2 data gradient and BN1 backward pass 1 are one
  // kernel per (image, M-tile) (bwd_mid_tc.cuh); slots 8 and 12 stay empty, slot 9 is the fused kernel
  bool mid = false;
  if constexpr (!std::is_same<T, float>::value) mid = c->use_tc && c->use_wgrad_tc && c->fuse_mid && fuse_bn2r && B <= 32;
  if (mid) {
    skip_slot();
  } else if (planes) {
    if constexpr (!std::is_same<T, float>::value) launch_k(k_bn2_bwd_apply_planes<T>, dim3(4, ns2p), dim3(256), 0, s, r2p, pdl);
  } else {
    launch_k(k_bn2_bwd_apply<T>, dim3(32, ns2), dim3(256), 0, s, r2, pdl);
  }
  if (!mid) mark();
  // 9. conv2 weight gradient partials, 10. conv2 data gradient
  const int gw = std::min(B, 37);
  Conv2WgradArgs w2{c->p1, c->dy2, c->gpart, B};
  if (fork && !mid) cudaEventRecord(c->ev_f2, s);  // the weight-gradient branch forks here, but is enqueued after the
  // data-gradient kernel below so that the critical-path kernel is dispatched first
  Conv14Args d2{c->dy2, c->params, c->dp1, nullptr, 0, st2, c->ds, c->comm, B, 0, 0, cfg.bn_eps};
  tc = false;
  if constexpr (!std::is_same<T, float>::value) {
    if (mid) {
      tc = true;
      // The conv2 weight-gradient branch forks BEFORE this kernel: it rebuilds d(conv2 out) itself (k_conv2_wgrad_tc<T, true>)
      // and only polls the BN2 backward sums this kernel's first CTA publishes, so it runs beside the backward middle
      // instead of behind it.  d(conv2 out) is therefore never written to HBM in this mode.
      if (fork) cudaEventRecord(c->ev_f2, s);
      BnBwdArgs r2m = r2;
      r2m.dy = nullptr;
      r2m.dy_planes = nullptr;
      BwdMidArgs ma{r2m, static_cast<char*>(c->w2_tiles) + TC_WBYTES, c->dp1, c->y1, st1, c->bpart1_ll};
      launch_k(k_bwd_mid_tc<T>, dim3(B, 2), dim3(MID_THREADS), MID_SMEM_BYTES, s, ma, pdl);
      if (fork) cudaEventRecord(c->ev_m, s);  // the side branch's row reduction also sums this kernel's conv2-bias rows
    } else if (c->use_tc) {
      tc = true;
      Conv2TcArgs t2{static_cast<char*>(c->w2_tiles) + TC_WBYTES, c->dy2h, c->params, c->dp1, nullptr, 0, st2, c->ds, c->comm, B, 0, 0, cfg.bn_eps};
      launch_k(k_conv2_tc<T, 2>, dim3(std::min(B, 296), tsplit), dim3(TC_THREADS), tc_smem_bytes(4), s, t2, pdl);
    }
  }
  if (!tc) k_conv14<T, 32, 16, 2><<<dim3(std::min(B, 296), 2), 128, (32 * 18 * C14_RS + 8 * 32 * 25) * 4, s>>>(d2);
  mark();
  if (fork) cudaStreamWaitEvent(s2, c->ev_f2, 0);
  int gw_rows = gw;
  bool wtc = false;
  if constexpr (!std::is_same<T, float>::value) {
    if (c->use_tc && c->use_wgrad_tc) {
      wtc = true;
      gw_rows = std::min(B, 148);
      // small batches: several images per CTA (the accumulators stay in TMEM) keep the 5-way kh split within one
      // wave of SMs and shorten the row reduction behind it; DDPSTEP_WG_DIV overrides the images-per-CTA choice
      static const int wg_div = [] { const char* e = getenv("DDPSTEP_WG_DIV"); return e ? std::max(1, atoi(e)) : 2; }();  // measured at batch 32: 1 -> 45.1, 2 -> 43.0, 4 -> 44.7 us per step
      if (B <= 64) gw_rows = cdiv(B, wg_div);
      Conv2WgradTcArgs wa{c->dy2h, c->p1h, c->gpart, B, r2};
      // small batches: split the 5 kh taps over 5 CTAs per image so more SMs take part
      if (mid) {
        gw_rows = B;  // one image per CTA: rebuilding d(conv2 out) is the long part of this kernel, so it is not serialised
        k_conv2_wgrad_tc<T, true><<<dim3(gw_rows, 5), WG_BUILD_THREADS, WG_SMEM_BYTES, s2>>>(wa);
      }
      else k_conv2_wgrad_tc<T, false><<<dim3(gw_rows, B <= 64 ? 5 : 1), TC_THREADS, WG_SMEM_BYTES, s2>>>(wa);
    }
  }
  if (!wtc) k_conv2_wgrad<T><<<dim3(gw, 4), 160, 0, s2>>>(w2);
  mark();
  if (fork && mid) cudaStreamWaitEvent(s2, c->ev_m, 0);
  // world > 1: the side branches also push their slices of the bucket to the peers (step_kernel.cuh, "early push")
  const int early = early_on ? 1 : 0;
  PushArgs pr{c->gpart, c->gred, gw_rows, mid ? 2 * B : 0, c->comm, c->ds, early, c->acc};
  k_reduce_rows<T><<<mid ? 26 : 25, 128, 0, s2>>>(pr); mark();
  c->last_push_c2 = pr;
  c->last_c2_ctas = mid ? 26 : 25;
  if (fork) cudaEventRecord(c->ev_j2, s2);
  // 11. BN1 backward reduce, 12. BN1 backward apply fused with conv1 weight gradient
  const int ns1 = mid ? 2 * B : clampi(B, 1, 32);  // one image per CTA at small batch: every thread has a single window, a single trip to L2
  BnBwdArgs r1{c->y1, c->dp1, c->params, st1, dstat1 == 2 ? c->bpart1_ll : c->bpart, c->gdirect, c->ds, c->comm, B, cdiv(B, ns1), 0, sync, dstat1, ns1, nullptr, nullptr, c->gpart};
  if (!mid) { launch_k(k_bn_bwd_reduce<T, 16, 28>, dim3(16, ns1), dim3(256), 0, s, r1, pdl); mark(); }
  else skip_slot();
  const int gc1 = std::min(B, 296);
  Conv1WgradArgs cw{r1, x};
  if (rec) rec->a_c1w = cw;
  int rows_c1 = gc1;
  bool c1wtc = false;
  if constexpr (!std::is_same<T, float>::value) {
    if (c1tc && c->use_conv1_wgrad_tc) {  // large batch: tensor-core weight gradient (conv1_tc.cuh)
      c1wtc = true;
      rows_c1 = std::min(B, 296);  // two CTAs per SM: the builders are latency-bound, a second CTA hides it
      Conv1WgradTcArgs cwt{r1, x};
      if (rec) { rec->a_c1wtc = cwt; rec->c1wtc = true; }
      launch_k(k_conv1_wgrad_tc<T>, dim3(rows_c1), dim3(C1W_TC_THREADS), C1W_TC_SMEM, s, cwt, pdl);
    }
  }
  if (!c1wtc) launch_k(k_conv1_wgrad<T>, dim3(gc1, 4), dim3(C1W_THREADS), 0, s, cw, pdl);
  mark();
  // 13. Reducer all-reduce + GradScaler + SGD
  StepArgs sa{c->params, c->grads, c->grads_scaled, c->gpart, c->gred, c->gdirect, c->nll, c->host_mail, c->bnbuf, c->nbt, c->ds, c->comm, B,
              rows_c1, gw_rows, mid ? 0 : (planes ? ns2p : ns2), nsf, cfg.use_scaler, sync, cfg.growth_factor, cfg.backoff_factor, cfg.growth_interval,
              c->mom, c->post_loss ? 1 : 0, c->acc, c->accumulate ? 1 : 0, c->tail_flags, early};
  if (!c->accumulate) { c->last_step = sa; c->last_step_dtype_ok = 1; }
  if (fork) { cudaStreamWaitEvent(s, c->ev_j1, 0); cudaStreamWaitEvent(s, c->ev_j2, 0); }
  if (c->tail_grid) launch_k(k_step_grid<T>, dim3(STEP_GRID_CTAS), dim3(STEP_GRID_THREADS), 0, s, sa, pdl);
  else launch_k(k_step<T>, dim3(STEP_CTAS), dim3(STEP_THREADS), 0, s, sa, pdl);
  mark();
  return cudaGetLastError();
}

template <typename T>
static cudaError_t enqueue_eval(ddpstep_ctx* c, const BatchSrc& x, const int64_t* y, int B, cudaStream_t s, Recorder* rec) {
  const ddpstep_config& cfg = c->cfg;
  const int dbl = cfg.conv_bias_double_round;
  auto mark = [&]() {
    if (rec) {
      rec->n++;
      if (rec->events) {
        cudaEvent_t ev;
        cudaEventCreate(&ev);
        if (rec->ev_external) cudaEventRecordWithFlags(ev, rec->ev_stream, cudaEventRecordExternal);
        else cudaEventRecord(ev, rec->ev_stream);
        rec->events->push_back(ev);
      }
    }
    c->launches++;
  };
  if constexpr (!std::is_same<T, float>::value) {
    if (c->use_tc) { k_refresh_w2_tiles<T><<<100, 256, 0, s>>>(c->params, c->w2_tiles); mark(); }
  }
  Conv1Args a1{x, c->params, c->y1, nullptr, 0, c->st, c->ds, c->comm, B, dbl, 0, cfg.bn_eps, nullptr, std::min(B, 296)};
  if (rec) rec->a_c1 = a1;
  k_conv1_fwd<T><<<dim3(std::min(B, 296), C1_GROUPS), C1_THREADS, 0, s>>>(a1); mark();
  BnPoolArgs b1{c->y1, c->p1, c->use_tc ? c->p1h : nullptr, c->params, c->st, c->bnbuf, c->nbt, c->ds, c->comm, B, 0, 1, 0, cfg.bn_momentum, cfg.bn_eps, nullptr, 0};
  bool planes = false;
  if constexpr (!std::is_same<T, float>::value) planes = c->use_tc && B > 64;
  c->last_planes = planes;
  if (planes) {
    if constexpr (!std::is_same<T, float>::value) k_bn1_relu_pool_planes<T><<<clampi(cdiv(B * 2 * 196, 256), 1, 1184), 256, 0, s>>>(b1);
  } else {
    k_bn_relu_pool<T, 16, 28><<<clampi(cdiv(B * 16 * 196, 256), 1, 1184), 256, 0, s>>>(b1);
  }
  mark();
  Conv14Args a2{c->p1, c->params, c->y2, nullptr, 0, c->st + 1, c->ds, c->comm, B, dbl, 0, cfg.bn_eps};
  bool tc = false;
  if constexpr (!std::is_same<T, float>::value) {
    if (c->use_tc) {
      tc = true;
      Conv2TcArgs t2{c->w2_tiles, c->p1h, c->params, c->y2, nullptr, 0, c->st + 1, c->ds, c->comm, B, dbl, 0, cfg.bn_eps};
      k_conv2_tc<T, 0><<<std::min(B, 148), TC_THREADS, tc_smem_bytes(2), s>>>(t2);
    }
  }
  if (!tc) k_conv14<T, 16, 32, 0><<<dim3(std::min(B, 148), 4), 128, (16 * 18 * C14_RS + 8 * 16 * 25) * 4, s>>>(a2);
  mark();
  BnPoolArgs b2{c->y2, c->p2, nullptr, c->params, c->st + 1, c->bnbuf, c->nbt, c->ds, c->comm, B, 1, 1, 0, cfg.bn_momentum, cfg.bn_eps, nullptr, 0};
  k_bn_relu_pool<T, 32, 14><<<clampi(cdiv(B * 32 * 49, 256), 1, 1184), 256, 0, s>>>(b2); mark();
  FcArgs f{c->p2, c->params, reinterpret_cast<const long long*>(y), x.idx, c->logits, nullptr, nullptr, nullptr, c->counts, c->ds, B, 0, 1,
           nullptr, nullptr, nullptr, BnPoolArgs{}, 0};
  if (rec) rec->a_fc = f;
  if (B > 1184) k_fc_ce<T, 4><<<296, 224, 0, s>>>(f);
  else k_fc_ce<T, 1><<<std::min(B, 296), 224, 0, s>>>(f);
  mark();
  return cudaGetLastError();
}

template <bool TRAIN>
static cudaError_t enqueue(ddpstep_ctx* c, const BatchSrc& x, const int64_t* y, int B, cudaStream_t s, Recorder* rec) {
  switch (c->cfg.dtype) {
    case DDPSTEP_FP32: return TRAIN ? enqueue_train<float>(c, x, y, B, s, rec) : enqueue_eval<float>(c, x, y, B, s, rec);
    case DDPSTEP_FP16: return TRAIN ? enqueue_train<__half>(c, x, y, B, s, rec) : enqueue_eval<__half>(c, x, y, B, s, rec);
    default: return TRAIN ? enqueue_train<__nv_bfloat16>(c, x, y, B, s, rec) : enqueue_eval<__nv_bfloat16>(c, x, y, B, s, rec);
  }
}

struct KernelPtrs { const void *c1, *fc, *c1w, *c1tc, *c1wtc, *fc4; };
static KernelPtrs kernel_ptrs(int dtype) {
  switch (dtype) {
    case DDPSTEP_FP32: return {(const void*)k_conv1_fwd<float>, (const void*)k_fc_ce<float, 1>, (const void*)k_conv1_wgrad<float>, nullptr, nullptr, (const void*)k_fc_ce<float, 4>};
    case DDPSTEP_FP16: return {(const void*)k_conv1_fwd<__half>, (const void*)k_fc_ce<__half, 1>, (const void*)k_conv1_wgrad<__half>, (const void*)k_conv1_tc<__half>, (const void*)k_conv1_wgrad_tc<__half>, (const void*)k_fc_ce<__half, 4>};
    default: return {(const void*)k_conv1_fwd<__nv_bfloat16>, (const void*)k_fc_ce<__nv_bfloat16, 1>, (const void*)k_conv1_wgrad<__nv_bfloat16>, (const void*)k_conv1_tc<__nv_bfloat16>, (const void*)k_conv1_wgrad_tc<__nv_bfloat16>, (const void*)k_fc_ce<__nv_bfloat16, 4>};
  }
}

static int repoint(ddpstep_ctx* ctx, GraphEntry& ge, cudaGraphNode_t node, void* arg_struct) {
  cudaKernelNodeParams kp;
  CK(cudaGraphKernelNodeGetParams(node, &kp));
  void* args[1] = {arg_struct};
  kp.kernelParams = args;
  kp.extra = nullptr;
  CK(cudaGraphExecKernelNodeSetParams(ge.exec, node, &kp));
  return 0;
}

// One captured graph per (mode, batch).  The step is launch- and latency-bound at batch 32,
// so the 13 kernels are replayed with a single cudaGraphLaunch.
template <bool TRAIN>
static int run_graph(ddpstep_ctx* ctx, const BatchSrc& x, const int64_t* y, int B, cudaStream_t s) {
  const GraphKey key(TRAIN ? (ctx->accumulate ? 2 : 0) : 1, x.x == nullptr ? 1 : 0, y != nullptr ? 1 : 0, B);
  auto it = ctx->graphs.find(key);
  if (it == ctx->graphs.end()) {
    GraphEntry ge;
    Recorder rec;
    const int64_t before = ctx->launches;
    CK(cudaStreamBeginCapture(ctx->cap_stream, cudaStreamCaptureModeThreadLocal));
    cudaError_t e = enqueue<TRAIN>(ctx, x, y, B, ctx->cap_stream, &rec);
    cudaError_t e2 = cudaStreamEndCapture(ctx->cap_stream, &ge.graph);
    ctx->launches = before;
    if (e != cudaSuccess || e2 != cudaSuccess) {
      if (ge.graph) cudaGraphDestroy(ge.graph);
      return fail(ctx, DDPSTEP_E_CUDA, "graph capture failed: %s", cudaGetErrorString(e != cudaSuccess ? e : e2));
    }
    if (cudaError_t e3 = cudaGraphInstantiate(&ge.exec, ge.graph, 0); e3 != cudaSuccess) {
      cudaGraphDestroy(ge.graph);
      return fail(ctx, DDPSTEP_E_CUDA, "cudaGraphInstantiate failed: %s", cudaGetErrorString(e3));
    }
    size_t n = 0;
    CK(cudaGraphGetNodes(ge.graph, nullptr, &n));
    std::vector<cudaGraphNode_t> nodes(n);
    CK(cudaGraphGetNodes(ge.graph, nodes.data(), &n));
    const KernelPtrs kp = kernel_ptrs(ctx->cfg.dtype);
    for (cudaGraphNode_t nd : nodes) {
      cudaGraphNodeType ty;
      CK(cudaGraphNodeGetType(nd, &ty));
      if (ty != cudaGraphNodeTypeKernel) continue;
      cudaKernelNodeParams p;
      CK(cudaGraphKernelNodeGetParams(nd, &p));
      if (p.func == kp.c1) ge.node_c1 = nd;
      else if (kp.c1tc && p.func == kp.c1tc) { ge.node_c1 = nd; ge.c1tc = true; }
      else if (p.func == kp.fc || p.func == kp.fc4) ge.node_fc = nd;
      else if (p.func == kp.c1w) ge.node_c1w = nd;
      else if (kp.c1wtc && p.func == kp.c1wtc) { ge.node_c1w = nd; ge.c1wtc = true; }
    }
    if (!ge.node_c1 || !ge.node_fc || (TRAIN && !ge.node_c1w)) {
      cudaGraphExecDestroy(ge.exec); cudaGraphDestroy(ge.graph);
      return fail(ctx, DDPSTEP_E_STATE, "could not locate the input-consuming kernel nodes in the captured graph");
    }
    ge.a_c1 = rec.a_c1; ge.a_fc = rec.a_fc; ge.a_c1w = rec.a_c1w; ge.a_c1tc = rec.a_c1tc; ge.a_c1wtc = rec.a_c1wtc;
    ge.x = x; ge.y = y; ge.n_kernels = rec.n;
    ge.planes = ctx->last_planes;  // set by enqueue_* during the capture
    it = ctx->graphs.emplace(key, ge).first;
  }
  GraphEntry& ge = it->second;
  if (ge.x.x != x.x || ge.x.u8 != x.u8 || ge.x.idx != x.idx) {
    ge.a_c1.x = x;
    ge.a_c1tc.x = x;
    int rc = ge.c1tc ? repoint(ctx, ge, ge.node_c1, &ge.a_c1tc) : repoint(ctx, ge, ge.node_c1, &ge.a_c1);
    if (rc) return rc;
    if (TRAIN) {
      ge.a_c1w.x = x;
      ge.a_c1wtc.x = x;
      rc = ge.c1wtc ? repoint(ctx, ge, ge.node_c1w, &ge.a_c1wtc) : repoint(ctx, ge, ge.node_c1w, &ge.a_c1w);
      if (rc) return rc;
    }
    ge.x = x;
  }
  if (ge.y != y || ge.a_fc.idx != x.idx) {
    ge.a_fc.y = reinterpret_cast<const long long*>(y);
    ge.a_fc.idx = x.idx;
    int rc = repoint(ctx, ge, ge.node_fc, &ge.a_fc);
    if (rc) return rc;
    ge.y = y;
  }
  CK(cudaGraphLaunch(ge.exec, s));
  ctx->launches += ge.n_kernels;
  ctx->last_planes = ge.planes;
  return 0;
}

static bool is_pinned_host(const void* p) {
  cudaPointerAttributes at;
  if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return false; }
  return at.type == cudaMemoryTypeHost;
}

// The training step fed from pinned host memory as ONE graph: the two H2D copies are root nodes (the label copy on
// a parallel branch), conv1 depends on both, the rest is the usual step.  Against two cudaMemcpyAsync calls plus a
// graph launch this saves two driver calls on the host and the stream-order gaps between three separately
// submitted pieces of work.  A new pair of host pointers re-points the two copy nodes of the instantiated graph.
static int run_graph_host(ddpstep_ctx* ctx, const float* xh, const int64_t* yh, int B, cudaStream_t s) {
  const GraphKey key(0, 2, 1, B);
  const size_t xb = (size_t)B * 784 * 4, yb = (size_t)B * 8;
  auto it = ctx->graphs.find(key);
  if (it == ctx->graphs.end()) {
    GraphEntry ge;
    Recorder rec;
    const int64_t before = ctx->launches;
    CK(cudaStreamBeginCapture(ctx->cap_stream, cudaStreamCaptureModeThreadLocal));
    cudaEventRecord(ctx->ev_h0, ctx->cap_stream);
    cudaStreamWaitEvent(ctx->side1, ctx->ev_h0, 0);
    cudaMemcpyAsync(ctx->y_stage, yh, yb, cudaMemcpyHostToDevice, ctx->side1);
    cudaEventRecord(ctx->ev_h1, ctx->side1);
    cudaMemcpyAsync(ctx->x_stage, xh, xb, cudaMemcpyHostToDevice, ctx->cap_stream);
    cudaStreamWaitEvent(ctx->cap_stream, ctx->ev_h1, 0);
    ctx->post_loss = true;
    cudaError_t e = enqueue<true>(ctx, BatchSrc{ctx->x_stage, nullptr, nullptr}, ctx->y_stage, B, ctx->cap_stream, &rec);
    ctx->post_loss = false;
    cudaError_t e2 = cudaStreamEndCapture(ctx->cap_stream, &ge.graph);
    ctx->launches = before;
    if (e != cudaSuccess || e2 != cudaSuccess) {
      if (ge.graph) cudaGraphDestroy(ge.graph);
      return fail(ctx, DDPSTEP_E_CUDA, "graph capture failed: %s", cudaGetErrorString(e != cudaSuccess ? e : e2));
    }
    if (cudaError_t e3 = cudaGraphInstantiate(&ge.exec, ge.graph, 0); e3 != cudaSuccess) {
      cudaGraphDestroy(ge.graph);
      return fail(ctx, DDPSTEP_E_CUDA, "cudaGraphInstantiate failed: %s", cudaGetErrorString(e3));
    }
    size_t n = 0;
    CK(cudaGraphGetNodes(ge.graph, nullptr, &n));
    std::vector<cudaGraphNode_t> nodes(n);
    CK(cudaGraphGetNodes(ge.graph, nodes.data(), &n));
    for (cudaGraphNode_t nd : nodes) {
      cudaGraphNodeType ty;
      CK(cudaGraphNodeGetType(nd, &ty));
      if (ty != cudaGraphNodeTypeMemcpy) continue;
      cudaMemcpy3DParms mp;
      CK(cudaGraphMemcpyNodeGetParams(nd, &mp));
      if (mp.dstPtr.ptr == ctx->x_stage) ge.node_hx = nd;
      else if (mp.dstPtr.ptr == ctx->y_stage) ge.node_hy = nd;
    }
    if (!ge.node_hx || !ge.node_hy) {
      cudaGraphExecDestroy(ge.exec); cudaGraphDestroy(ge.graph);
      return fail(ctx, DDPSTEP_E_STATE, "could not locate the copy nodes in the captured graph");
    }
    ge.hx = xh; ge.hy = yh; ge.n_kernels = rec.n;
    ge.planes = ctx->last_planes;
    it = ctx->graphs.emplace(key, ge).first;
  }
  GraphEntry& ge = it->second;
  if (ge.hx != xh) {
    CK(cudaGraphExecMemcpyNodeSetParams1D(ge.exec, ge.node_hx, ctx->x_stage, xh, xb, cudaMemcpyHostToDevice));
    ge.hx = xh;
  }
  if (ge.hy != yh) {
    CK(cudaGraphExecMemcpyNodeSetParams1D(ge.exec, ge.node_hy, ctx->y_stage, yh, yb, cudaMemcpyHostToDevice));
    ge.hy = yh;
  }
  CK(cudaGraphLaunch(ge.exec, s));
  ctx->launches += ge.n_kernels;
  ctx->last_planes = ge.planes;
  return 0;
}

// A cross-rank wait timed out in an earlier step: the tail kernel raised word 2 of the pinned mailbox.  From then on
// every step entry point refuses to launch (the replicas can no longer be assumed identical).
static int peer_failed(ddpstep_ctx* ctx) {
  if (ctx->host_mail && reinterpret_cast<const volatile int*>(ctx->host_mail)[2] != 0)
    return fail(ctx, DDPSTEP_E_PEER, "a peer did not answer within the timeout in an earlier step; that step and all later "
                                     "ones applied no update -- this context is dead (restart the job from a checkpoint)");
  return 0;
}

static int check_step_args(ddpstep_ctx* ctx, const void* x, int32_t batch) {
  if (!ctx) return DDPSTEP_E_ARG;
  if (!x) return fail(ctx, DDPSTEP_E_ARG, "null input pointer");
  if (reinterpret_cast<uintptr_t>(x) & 15) return fail(ctx, DDPSTEP_E_ARG, "the image pointer must be 16-byte aligned");
  if (batch < 1 || batch > ctx->cfg.batch_cap)
    return fail(ctx, DDPSTEP_E_ARG, "batch %d outside [1, batch_cap=%d]", batch, ctx->cfg.batch_cap);
  if (!ctx->peers_ready) return fail(ctx, DDPSTEP_E_STATE, "world > 1: call ddpstep_import_peers first");
  return peer_failed(ctx);
}

extern "C" int ddpstep_train_step(ddpstep_ctx* ctx, const float* x, const int64_t* y, int32_t batch, void* stream) {
  int rc = check_step_args(ctx, x, batch);
  if (rc) return rc;
  if (!y) return fail(ctx, DDPSTEP_E_ARG, "null label pointer");
  CK(cudaSetDevice(ctx->cfg.device));
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const BatchSrc src{x, nullptr, nullptr};
  if (ctx->cfg.use_graph) {
    rc = run_graph<true>(ctx, src, y, batch, s);
  } else {
    cudaError_t e = enqueue<true>(ctx, src, y, batch, s, nullptr);
    if (e != cudaSuccess) rc = fail(ctx, DDPSTEP_E_CUDA, "train step launch failed: %s", cudaGetErrorString(e));
  }
  if (rc == 0) ctx->steps_enqueued++;  // only a step that was really launched will post its sequence number
  return rc;
}

// K consecutive iterations of train() (ddp_main.py:85-93) as ONE graph launch: the steps are chained inside the graph by
// programmatic edges (the next step's conv1 is resident and waiting while the tail kernel of the current one runs), so
// the ~3 us between two graph launches and the host's launch jitter -- at world > 1 the first exchange of every step
// absorbs the slowest rank's -- are paid once per K steps instead of once per step.  The batches are copied (device to
// device, stream-ordered) into staging buffers of the library, so the captured graph is replayed for any pointers.
//   kind 0: x_dev [K*batch, 784] fp32 + y_dev [K*batch] int64, K consecutive batches
//   kind 1: uint8 dataset + labels + idx_dev [K*batch] sampler indices (a whole epoch segment per call)
static int run_multi(ddpstep_ctx* ctx, int kind, const void* data, const int64_t* labels, const int64_t* idx, int K, int B,
                     cudaStream_t s) {
  if (K < 1 || K > 4096) return fail(ctx, DDPSTEP_E_ARG, "n_steps %d outside [1, 4096]", K);
  if (!ctx->cfg.use_graph) return fail(ctx, DDPSTEP_E_STATE, "multi-step launches need use_graph = 1");
  const size_t need = (size_t)K * B;
  if (need > ctx->multi_cap) {  // lazily sized staging: allocation happens outside any capture, old graphs are dropped
    CK(cudaStreamSynchronize(s));
    for (auto& kv : ctx->multi_graphs) { cudaGraphExecDestroy(kv.second.exec); cudaGraphDestroy(kv.second.graph); }
    ctx->multi_graphs.clear();
    cudaFree(ctx->xs_stage); cudaFree(ctx->ys_stage); cudaFree(ctx->idx_stage);
    ctx->xs_stage = nullptr; ctx->ys_stage = nullptr; ctx->idx_stage = nullptr; ctx->multi_cap = 0;
    CK(cudaMalloc(&ctx->xs_stage, need * 784 * 4));
    CK(cudaMalloc(&ctx->ys_stage, need * 8));
    CK(cudaMalloc(&ctx->idx_stage, need * 8));
    ctx->multi_cap = need;
  }
  if (kind == 0) {
    CK(cudaMemcpyAsync(ctx->xs_stage, data, need * 784 * 4, cudaMemcpyDeviceToDevice, s));
    CK(cudaMemcpyAsync(ctx->ys_stage, labels, need * 8, cudaMemcpyDeviceToDevice, s));
  } else {
    CK(cudaMemcpyAsync(ctx->idx_stage, idx, need * 8, cudaMemcpyDeviceToDevice, s));
  }
  const auto key = std::make_tuple(kind, K, B, kind == 0 ? (const void*)nullptr : data, kind == 0 ? (const void*)nullptr : (const void*)labels);
  auto it = ctx->multi_graphs.find(key);
  if (it == ctx->multi_graphs.end()) {
    GraphEntry ge;
    Recorder rec;
    const int64_t before = ctx->launches;
    CK(cudaStreamBeginCapture(ctx->cap_stream, cudaStreamCaptureModeThreadLocal));
    cudaError_t e = cudaSuccess;
    for (int k = 0; k < K && e == cudaSuccess; ++k) {
      ctx->chained_step = k > 0;
      const BatchSrc src = kind == 0 ? BatchSrc{ctx->xs_stage + (size_t)k * B * 784, nullptr, nullptr}
                                     : BatchSrc{nullptr, static_cast<const unsigned char*>(data), ctx->idx_stage + (size_t)k * B};
      e = enqueue<true>(ctx, src, kind == 0 ? ctx->ys_stage + (size_t)k * B : labels, B, ctx->cap_stream, &rec);
    }
    ctx->chained_step = false;
    cudaError_t e2 = cudaStreamEndCapture(ctx->cap_stream, &ge.graph);
    ctx->launches = before;
    if (e != cudaSuccess || e2 != cudaSuccess) {
      if (ge.graph) cudaGraphDestroy(ge.graph);
      return fail(ctx, DDPSTEP_E_CUDA, "multi-step graph capture failed: %s", cudaGetErrorString(e != cudaSuccess ? e : e2));
    }
    if (cudaError_t e3 = cudaGraphInstantiate(&ge.exec, ge.graph, 0); e3 != cudaSuccess) {
      cudaGraphDestroy(ge.graph);
      return fail(ctx, DDPSTEP_E_CUDA, "cudaGraphInstantiate failed: %s", cudaGetErrorString(e3));
    }
    ge.n_kernels = rec.n;
    ge.planes = ctx->last_planes;
    it = ctx->multi_graphs.emplace(key, ge).first;
  }
  CK(cudaGraphLaunch(it->second.exec, s));
  ctx->launches += it->second.n_kernels;
  ctx->last_planes = it->second.planes;
  ctx->steps_enqueued += (unsigned int)K;
  return 0;
}

extern "C" int ddpstep_train_steps(ddpstep_ctx* ctx, const float* x_dev, const int64_t* y_dev, int32_t n_steps, int32_t batch,
                                   void* stream) {
  int rc = check_step_args(ctx, x_dev, batch);
  if (rc) return rc;
  if (!y_dev) return fail(ctx, DDPSTEP_E_ARG, "null label pointer");
  CK(cudaSetDevice(ctx->cfg.device));
  return run_multi(ctx, 0, x_dev, y_dev, nullptr, n_steps, batch, static_cast<cudaStream_t>(stream));
}

// A micro-step inside `with model.no_sync():` (torch/nn/parallel/distributed.py:1441-1470): forward, loss and
// backward exactly as ddpstep_train_step -- BN statistics, running buffers and, under SyncBN, their exchanges included --
// but the local gradient is ADDED to an accumulation buffer (what `.grad` does when zero_grad() is not called) and nothing
// is exchanged, unscaled or applied.  The next ddpstep_train_step* call all-reduces buffer + its own gradient.
extern "C" int ddpstep_train_step_accumulate(ddpstep_ctx* ctx, const float* x, const int64_t* y, int32_t batch, void* stream) {
  int rc = check_step_args(ctx, x, batch);
  if (rc) return rc;
  if (!y) return fail(ctx, DDPSTEP_E_ARG, "null label pointer");
  CK(cudaSetDevice(ctx->cfg.device));
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const BatchSrc src{x, nullptr, nullptr};
  ctx->accumulate = true;
  if (ctx->cfg.use_graph) {
    rc = run_graph<true>(ctx, src, y, batch, s);
  } else {
    cudaError_t e = enqueue<true>(ctx, src, y, batch, s, nullptr);
    if (e != cudaSuccess) rc = fail(ctx, DDPSTEP_E_CUDA, "train step launch failed: %s", cudaGetErrorString(e));
  }
  ctx->accumulate = false;
  if (rc == 0) ctx->steps_enqueued++;
  return rc;
}

// SURVEY 8(f) N1, the step immediately before the hot path: torchvision MNIST -> ToTensor -> DataLoader -> .cuda()
// (ddp_main.py:127-141,86-87) becomes: the uint8 dataset resident in HBM, the sampler's indices for this batch, and the
// gather + `/ 255` fused into conv1's load.
static int check_indexed(ddpstep_ctx* ctx, const void* images, const void* labels, const void* idx, int32_t batch) {
  if (!ctx) return DDPSTEP_E_ARG;
  if (!images || !idx) return fail(ctx, DDPSTEP_E_ARG, "null dataset / index pointer");
  if (reinterpret_cast<uintptr_t>(images) & 15) return fail(ctx, DDPSTEP_E_ARG, "the uint8 dataset must be 16-byte aligned");
  if (batch < 1 || batch > ctx->cfg.batch_cap)
    return fail(ctx, DDPSTEP_E_ARG, "batch %d outside [1, batch_cap=%d]", batch, ctx->cfg.batch_cap);
  if (!ctx->peers_ready) return fail(ctx, DDPSTEP_E_STATE, "world > 1: call ddpstep_import_peers first");
  (void)labels;
  return peer_failed(ctx);
}

extern "C" int ddpstep_train_step_indexed(ddpstep_ctx* ctx, const uint8_t* images_u8, const int64_t* labels,
                                          const int64_t* idx, int32_t batch, void* stream) {
  int rc = check_indexed(ctx, images_u8, labels, idx, batch);
  if (rc) return rc;
  if (!labels) return fail(ctx, DDPSTEP_E_ARG, "null label pointer");
  CK(cudaSetDevice(ctx->cfg.device));
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const BatchSrc src{nullptr, images_u8, reinterpret_cast<const long long*>(idx)};
  if (ctx->cfg.use_graph) {
    rc = run_graph<true>(ctx, src, labels, batch, s);
  } else {
    cudaError_t e = enqueue<true>(ctx, src, labels, batch, s, nullptr);
    if (e != cudaSuccess) rc = fail(ctx, DDPSTEP_E_CUDA, "train step launch failed: %s", cudaGetErrorString(e));
  }
  if (rc == 0) ctx->steps_enqueued++;
  return rc;
}

extern "C" int ddpstep_eval_batch_indexed(ddpstep_ctx* ctx, const uint8_t* images_u8, const int64_t* labels,
                                          const int64_t* idx, int32_t batch, void* stream) {
  int rc = check_indexed(ctx, images_u8, labels, idx, batch);
  if (rc) return rc;
  CK(cudaSetDevice(ctx->cfg.device));
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const BatchSrc src{nullptr, images_u8, reinterpret_cast<const long long*>(idx)};
  if (ctx->cfg.use_graph) return run_graph<false>(ctx, src, labels, batch, s);
  cudaError_t e = enqueue<false>(ctx, src, labels, batch, s, nullptr);
  if (e != cudaSuccess) return fail(ctx, DDPSTEP_E_CUDA, "eval launch failed: %s", cudaGetErrorString(e));
  return 0;
}

extern "C" int ddpstep_train_steps_indexed(ddpstep_ctx* ctx, const uint8_t* images_u8, const int64_t* labels,
                                           const int64_t* idx, int32_t n_steps, int32_t batch, void* stream) {
  int rc = check_indexed(ctx, images_u8, labels, idx, batch);
  if (rc) return rc;
  if (!labels) return fail(ctx, DDPSTEP_E_ARG, "null label pointer");
  CK(cudaSetDevice(ctx->cfg.device));
  return run_multi(ctx, 1, images_u8, labels, idx, n_steps, batch, static_cast<cudaStream_t>(stream));
}

// Is this host pointer page-locked and mapped (cudaHostAlloc / cudaHostRegister / torch pin_memory)?  Then
// the kernels can read it in place over PCIe and the two staging copies of ddp_main.py:86-87 disappear.
static const void* mapped_device_pointer(const void* host) {
  cudaPointerAttributes at;
  if (cudaPointerGetAttributes(&at, host) != cudaSuccess) { cudaGetLastError(); return nullptr; }
  if (at.type != cudaMemoryTypeHost || at.devicePointer == nullptr) return nullptr;
  return at.devicePointer;
}

extern "C" int ddpstep_train_step_host(ddpstep_ctx* ctx, const float* xh, const int64_t* yh, int32_t batch,
                                       float* loss_out, void* stream) {
  int rc = check_step_args(ctx, xh, batch);
  if (rc) return rc;
  if (!yh) return fail(ctx, DDPSTEP_E_ARG, "null label pointer");
  CK(cudaSetDevice(ctx->cfg.device));
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  // Reading pinned batches in place (DDPSTEP_ZEROCOPY=1) saves the two staging copies but makes conv1 wait on PCIe
  // twice and re-points three graph nodes every step; measured slower (98 vs 93 us) at batch 32, so staging is
  // the default.
  static const bool zero_copy = [] { const char* e = getenv("DDPSTEP_ZEROCOPY"); return e && e[0] == '1'; }();
  const float* xd = zero_copy ? static_cast<const float*>(mapped_device_pointer(xh)) : nullptr;
  const int64_t* yd = zero_copy ? static_cast<const int64_t*>(mapped_device_pointer(yh)) : nullptr;
  // pinned batches: copies and step as one graph.  DDPSTEP_HOSTGRAPH=0 keeps the separate copies for comparison.
  static const bool host_graph = [] { const char* e = getenv("DDPSTEP_HOSTGRAPH"); return !(e && e[0] == '0'); }();
  if (!zero_copy && host_graph && ctx->cfg.use_graph && is_pinned_host(xh) && is_pinned_host(yh)) {
    rc = run_graph_host(ctx, xh, yh, batch, s);
    if (rc) return rc;
    ctx->steps_enqueued++;
    xd = nullptr;
  } else if (xd == nullptr || yd == nullptr) {  // stage through the library's device buffers (fixed graph pointers)
    CK(cudaMemcpyAsync(ctx->x_stage, xh, (size_t)batch * 784 * 4, cudaMemcpyHostToDevice, s));
    CK(cudaMemcpyAsync(ctx->y_stage, yh, (size_t)batch * 8, cudaMemcpyHostToDevice, s));
    xd = ctx->x_stage;
    yd = ctx->y_stage;
  }
  if (xd != nullptr) {  // pageable buffers, no graph, or zero-copy: the ordinary step, and a real synchronisation for the loss
    rc = ddpstep_train_step(ctx, xd, yd, batch, stream);
    if (rc) return rc;
    if (loss_out) {
      CK(cudaMemcpyAsync(loss_out, &ctx->ds->loss, sizeof(float), cudaMemcpyDeviceToHost, s));
      CK(cudaStreamSynchronize(s));
      return peer_failed(ctx);
    }
    return 0;
  }
  if (loss_out) {
    // the tail kernel posts {loss, step counter} into pinned host memory: poll it instead of copying back and
    // synchronising the stream; fall back to a real synchronisation if nothing shows up (device error)
    const volatile unsigned int* seq = reinterpret_cast<const volatile unsigned int*>(ctx->host_mail) + 1;
    const unsigned int want = ctx->steps_enqueued;
    bool seen = false;
    for (long spin = 0; spin < 200000000L; ++spin) {
      if (*seq == want) { seen = true; break; }
      if ((spin & 0xFFFFF) == 0xFFFFF && cudaStreamQuery(s) != cudaErrorNotReady) break;
    }
    if (!seen) {
      CK(cudaStreamSynchronize(s));
      if (*seq != want) return fail(ctx, DDPSTEP_E_STATE, "the step finished without posting its loss");
    }
    *loss_out = ctx->host_mail[0];
    return peer_failed(ctx);  // the step just awaited may be the one that timed out
  }
  return 0;
}

extern "C" int ddpstep_eval_batch(ddpstep_ctx* ctx, const float* x, const int64_t* y, int32_t batch, void* stream) {
  int rc = check_step_args(ctx, x, batch);
  if (rc) return rc;
  CK(cudaSetDevice(ctx->cfg.device));
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const BatchSrc src{x, nullptr, nullptr};
  if (ctx->cfg.use_graph) return run_graph<false>(ctx, src, y, batch, s);
  cudaError_t e = enqueue<false>(ctx, src, y, batch, s, nullptr);
  if (e != cudaSuccess) return fail(ctx, DDPSTEP_E_CUDA, "eval launch failed: %s", cudaGetErrorString(e));
  return 0;
}

extern "C" int ddpstep_reset_counts(ddpstep_ctx* ctx, void* stream) {
  if (!ctx) return DDPSTEP_E_ARG;
  CK(cudaSetDevice(ctx->cfg.device));
  CK(cudaMemsetAsync(ctx->counts, 0, 8, static_cast<cudaStream_t>(stream)));
  return 0;
}

extern "C" int ddpstep_reduce_counts(ddpstep_ctx* ctx, void* stream) {
  if (!ctx) return DDPSTEP_E_ARG;
  if (!ctx->peers_ready) return fail(ctx, DDPSTEP_E_STATE, "world > 1: call ddpstep_import_peers first");
  CK(cudaSetDevice(ctx->cfg.device));
  CountsArgs a{ctx->counts, ctx->ds, ctx->comm};
  k_reduce_counts<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(a);
  ctx->launches++;
  CK(cudaGetLastError());
  return 0;
}

extern "C" int ddpstep_read_counts(ddpstep_ctx* ctx, float* correct, float* size, void* stream) {
  if (!ctx) return DDPSTEP_E_ARG;
  CK(cudaSetDevice(ctx->cfg.device));
  float h[2];
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  CK(cudaMemcpyAsync(h, ctx->counts, 8, cudaMemcpyDeviceToHost, s));
  CK(cudaStreamSynchronize(s));
  if (correct) *correct = h[0];
  if (size) *size = h[1];
  return 0;
}

extern "C" int ddpstep_read_activation(ddpstep_ctx* ctx, int32_t which, void* host, int64_t bytes) {
  if (!ctx || !host) return fail(ctx, DDPSTEP_E_ARG, "null argument");
  const size_t B = (size_t)ctx->cfg.batch_cap, es = (size_t)ctx->esize;
  const void* src[] = {ctx->y1, ctx->p1, ctx->y2, ctx->p2, ctx->dlogits, ctx->dy2, ctx->dp1};
  const size_t sz[] = {B * 16 * 784 * es, B * 16 * 196 * es, B * 32 * 196 * es, B * 1568 * es, B * 10 * es,
                       B * 32 * 196 * es, B * 16 * 196 * es};
  if (which < 0 || which > 6) return fail(ctx, DDPSTEP_E_ARG, "bad activation id %d", which);
  if (which == 5 && ctx->fuse_mid && ctx->use_tc)
    return fail(ctx, DDPSTEP_E_STATE, "d(conv2 out) is not materialised in HBM by the fused backward middle (DDPSTEP_MID=1)");
  CK(cudaSetDevice(ctx->cfg.device));
  CK(cudaDeviceSynchronize());
  const bool planes = ctx->last_planes && (which == 1 || which == 5);
  if (planes) {
    // on the tensor-core path p1 and d(conv2 out) exist only in the padded-plane layout: convert to NCHW here
    const int ch = which == 1 ? 16 : 32, npl = ch / 8;
    std::vector<uint16_t> pl((size_t)B * npl * 320 * 8);
    CK(cudaMemcpy(pl.data(), which == 1 ? ctx->p1h : ctx->dy2h, pl.size() * 2, cudaMemcpyDeviceToHost));
    uint16_t* out = static_cast<uint16_t*>(host);
    const size_t max_el = (size_t)bytes / 2;
    for (size_t n = 0; n < B; ++n)
      for (int c = 0; c < ch; ++c)
        for (int h = 0; h < 14; ++h)
          for (int w = 0; w < 14; ++w) {
            const size_t o = ((n * ch + c) * 14 + h) * 14 + w;
            if (o < max_el) out[o] = pl[((n * npl + (c >> 3)) * 320 + (h + 3) * 16 + w) * 8 + (c & 7)];
          }
    return 0;
  }
  CK(cudaMemcpy(host, src[which], std::min((size_t)bytes, sz[which]), cudaMemcpyDeviceToHost));
  return 0;
}

constexpr int kTrainKernels = 15;
static const char* const kTrainKernelNames[kTrainKernels] = {
    "w2_tiles_refresh", "conv1_fwd", "bn1_relu_pool", "conv2_fwd", "bn2_relu_pool", "fc_ce", "fc_wgrad", "bn2_bwd_reduce",
    "bn2_bwd_apply", "conv2_dgrad", "conv2_wgrad", "conv2_wgrad_reduce", "bn1_bwd_reduce", "conv1_wgrad", "step"};

extern "C" const char* ddpstep_kernel_name(int32_t i) { return (i >= 0 && i < kTrainKernels) ? kTrainKernelNames[i] : ""; }

extern "C" int ddpstep_profile_step(ddpstep_ctx* ctx, const float* x, const int64_t* y, int32_t batch, int32_t in_graph,
                                    float* ms_out, int32_t n_out, void* stream) {
  int rc = check_step_args(ctx, x, batch);
  if (rc) return rc;
  if (!y || !ms_out || n_out < kTrainKernels) return fail(ctx, DDPSTEP_E_ARG, "need labels and room for 15 durations");
  CK(cudaSetDevice(ctx->cfg.device));
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  std::vector<cudaEvent_t> evs;
  cudaEvent_t ev0;
  CK(cudaEventCreate(&ev0));
  Recorder rec;
  rec.events = &evs;
  cudaError_t e;
  auto drop_events = [&] { cudaEventDestroy(ev0); for (cudaEvent_t ev : evs) cudaEventDestroy(ev); };
  if (in_graph) {
    cudaGraph_t graph;
    cudaGraphExec_t exec;
    rec.ev_stream = ctx->cap_stream;
    rec.ev_external = true;
    const int64_t before = ctx->launches;
    CK(cudaStreamBeginCapture(ctx->cap_stream, cudaStreamCaptureModeThreadLocal));
    cudaEventRecordWithFlags(ev0, ctx->cap_stream, cudaEventRecordExternal);
    e = enqueue<true>(ctx, BatchSrc{x, nullptr, nullptr}, y, batch, ctx->cap_stream, &rec);
    cudaError_t e2 = cudaStreamEndCapture(ctx->cap_stream, &graph);
    ctx->launches = before;
    if (e != cudaSuccess || e2 != cudaSuccess) {
      if (e2 == cudaSuccess && graph) cudaGraphDestroy(graph);
      drop_events();
      return fail(ctx, DDPSTEP_E_CUDA, "profile capture failed: %s", cudaGetErrorString(e != cudaSuccess ? e : e2));
    }
    e = cudaGraphInstantiate(&exec, graph, 0);
    if (e == cudaSuccess) {
      e = cudaGraphLaunch(exec, s);
      if (e == cudaSuccess) { ctx->launches += rec.n; ctx->steps_enqueued++; e = cudaStreamSynchronize(s); }
      cudaGraphExecDestroy(exec);
    }
    cudaGraphDestroy(graph);
    if (e != cudaSuccess) { drop_events(); return fail(ctx, DDPSTEP_E_CUDA, "profile graph failed: %s", cudaGetErrorString(e)); }
  } else {
    rec.ev_stream = s;
    e = cudaEventRecord(ev0, s);
    if (e == cudaSuccess) e = enqueue<true>(ctx, BatchSrc{x, nullptr, nullptr}, y, batch, s, &rec);
    if (e == cudaSuccess) { ctx->steps_enqueued++; e = cudaStreamSynchronize(s); }
    if (e != cudaSuccess) { drop_events(); return fail(ctx, DDPSTEP_E_CUDA, "profile launch failed: %s", cudaGetErrorString(e)); }
  }
  cudaEvent_t prev = ev0;
  for (size_t i = 0; i < evs.size() && i < (size_t)kTrainKernels; ++i) {
    if (cudaEventElapsedTime(&ms_out[i], prev, evs[i]) != cudaSuccess) { cudaGetLastError(); ms_out[i] = 0.f; }
    if (rec.skipped >> i & 1u) ms_out[i] = -ms_out[i];  // nothing ran in this slot: minus the bare event-to-event interval
    prev = evs[i];
  }
  drop_events();
  return 0;
}

extern "C" int64_t ddpstep_launch_count(const ddpstep_ctx* ctx) { return ctx ? ctx->launches : 0; }

extern "C" int ddpstep_device_status(ddpstep_ctx* ctx) {
  if (!ctx) return DDPSTEP_E_ARG;
  CK(cudaSetDevice(ctx->cfg.device));
  CK(cudaDeviceSynchronize());
  DevState h;
  CK(cudaMemcpy(&h, ctx->ds, sizeof h, cudaMemcpyDeviceToHost));
  if (h.error) return fail(ctx, DDPSTEP_E_PEER, "a peer did not answer within the spin timeout; steps since then applied no update");
  return 0;
}

// torch.optim.SGD's param_groups[0] at run time (an LR scheduler, a hand-edited group): the values live in device
// memory, so neither the context nor any captured graph has to be rebuilt.  Stream-ordered: steps enqueued after this
// call on the same stream use the new values.  ddp_main.py:125 sets them once; torch/optim/sgd.py:425-465 reads them
// every step.
extern "C" int ddpstep_set_sgd_group(ddpstep_ctx* ctx, uint32_t tensor_mask, float lr, float momentum, float weight_decay,
                                     float dampening, int32_t nesterov, void* stream) {
  if (!ctx) return DDPSTEP_E_ARG;
  if ((tensor_mask & 0x3FFu) == 0 || (tensor_mask & ~0x3FFu))
    return fail(ctx, DDPSTEP_E_ARG, "tensor_mask must select some of the 10 parameter tensors (bits 0..9)");
  if (momentum < 0.f || weight_decay < 0.f) return fail(ctx, DDPSTEP_E_ARG, "momentum and weight_decay must be >= 0");
  if (nesterov && (momentum <= 0.f || dampening != 0.f))
    return fail(ctx, DDPSTEP_E_ARG, "nesterov momentum requires a momentum and zero dampening");  // sgd.py:68
  CK(cudaSetDevice(ctx->cfg.device));
  const DevState::SgdGroup g = {lr, momentum, weight_decay, dampening, nesterov ? 1 : 0};
  for (int t = 0; t < 10; ++t)
    if (tensor_mask >> t & 1u) {
      ctx->groups[t] = g;
      // pageable source: the runtime stages it before returning
      CK(cudaMemcpyAsync(&ctx->ds->grp[t], &ctx->groups[t], sizeof g, cudaMemcpyHostToDevice, static_cast<cudaStream_t>(stream)));
    }
  return 0;
}

extern "C" int ddpstep_set_sgd(ddpstep_ctx* ctx, float lr, float momentum, float weight_decay, float dampening, int32_t nesterov,
                               void* stream) {
  int rc = ddpstep_set_sgd_group(ctx, 0x3FFu, lr, momentum, weight_decay, dampening, nesterov, stream);
  if (rc == 0) {
    ctx->cfg.lr = lr; ctx->cfg.momentum = momentum; ctx->cfg.weight_decay = weight_decay; ctx->cfg.dampening = dampening;
    ctx->cfg.nesterov = nesterov ? 1 : 0;
  }
  return rc;
}

// DDP communication-hook work for one GradBucket of any model (SURVEY 8b alternative B, 8f N3): bucket <- sum over
// ranks of bucket * (1/W) when `average`, plain sum otherwise, in place, identical bits on every rank.  Buckets longer
// than 29,056 elements are reduced chunk by chunk through the same double-buffered inbox the step's own gradient uses.
// Collective and stream-ordered; must not be interleaved with an in-flight training step on another stream.
extern "C" int ddpstep_allreduce_bucket(ddpstep_ctx* ctx, float* bucket_dev, int64_t n, int32_t average, void* stream) {
  if (!ctx || !bucket_dev || n < 0) return fail(ctx, DDPSTEP_E_ARG, "null bucket / negative length");
  if (reinterpret_cast<uintptr_t>(bucket_dev) & 3) return fail(ctx, DDPSTEP_E_ARG, "the bucket must be 4-byte aligned");
  if (!ctx->peers_ready) return fail(ctx, DDPSTEP_E_STATE, "world > 1: call ddpstep_import_peers first");
  int rc = peer_failed(ctx);
  if (rc) return rc;
  CK(cudaSetDevice(ctx->cfg.device));
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const float pre = average ? 1.0f / (float)ctx->cfg.world : 1.0f;
  for (int64_t off = 0; off < n; off += N_PAD) {
    const int len = (int)std::min<int64_t>(N_PAD, n - off);
    BucketArgs ba{bucket_dev + off, len, pre, ctx->ds, ctx->comm};
    k_bucket_allreduce<<<cdiv(len, 1024), 256, 0, s>>>(ba);
    k_bump_sync_epoch<<<1, 1, 0, s>>>(ctx->ds);
    ctx->launches += 2;
  }
  CK(cudaGetLastError());
  return 0;
}

// How long a kernel polls for a peer's words before it gives up (default 20 s, or DDPSTEP_PEER_TIMEOUT_MS).  NCCL in the
// reference blocks until its watchdog fires; here a step whose wait expired applies no update and
// every later call returns DDPSTEP_E_PEER.
extern "C" int ddpstep_set_peer_timeout_ms(ddpstep_ctx* ctx, double ms) {
  if (!ctx || !(ms > 0.0)) return fail(ctx, DDPSTEP_E_ARG, "timeout must be positive");
  CK(cudaSetDevice(ctx->cfg.device));
  const long long cycles = (long long)(ms * (double)ctx->sm_khz);
  CK(cudaMemcpy(&ctx->ds->timeout_cycles, &cycles, sizeof cycles, cudaMemcpyHostToDevice));
  return 0;
}

// The tail kernel on its own: Reducer all-reduce of the 116,136-byte bucket fused with unscale / non-finite check /
// SGD / scale update, re-run on the partial gradients the last training step left behind (state advances like a
// step: the same gradient is applied again).  For measuring the exchange in isolation (tools/allreduce_bench.py,
// bench.py `allreduce`); every rank must call it the same number of times.  Replaces the single ncclAllReduce behind
// ddp_main.py:121-123 (reducer.hpp:524) + grad_scaler.py:294-298,371 + sgd.py:473.
extern "C" int ddpstep_allreduce_only(ddpstep_ctx* ctx, void* stream) {
  if (!ctx) return DDPSTEP_E_ARG;
  if (!ctx->last_step_dtype_ok) return fail(ctx, DDPSTEP_E_STATE, "run a training step first: the tail kernel re-reduces its partial gradients");
  int rc = peer_failed(ctx);
  if (rc) return rc;
  CK(cudaSetDevice(ctx->cfg.device));
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  StepArgs sa = ctx->last_step;
  sa.post_loss = 0;
  const int nfc = cdiv(N_PAD - OFF_WFC, 512), nc2 = ctx->last_c2_ctas;
  // with early push the exchange is three launches: the two side-branch kernels that finish + push 98 % of the bucket,
  // then the tail (in a step the first two run long before the tail, here back to back)
  switch (ctx->cfg.dtype) {
    case DDPSTEP_FP32:
      if (sa.early_push) { k_fc_reduce_push<float><<<nfc, 128, 0, s>>>(ctx->last_push_fc); k_reduce_rows<float><<<nc2, 128, 0, s>>>(ctx->last_push_c2); }
      if (ctx->tail_grid) launch_k(k_step_grid<float>, dim3(STEP_GRID_CTAS), dim3(STEP_GRID_THREADS), 0, s, sa, false);
      else launch_k(k_step<float>, dim3(STEP_CTAS), dim3(STEP_THREADS), 0, s, sa, false);
      break;
    case DDPSTEP_FP16:
      if (sa.early_push) { k_fc_reduce_push<__half><<<nfc, 128, 0, s>>>(ctx->last_push_fc); k_reduce_rows<__half><<<nc2, 128, 0, s>>>(ctx->last_push_c2); }
      if (ctx->tail_grid) launch_k(k_step_grid<__half>, dim3(STEP_GRID_CTAS), dim3(STEP_GRID_THREADS), 0, s, sa, false);
      else launch_k(k_step<__half>, dim3(STEP_CTAS), dim3(STEP_THREADS), 0, s, sa, false);
      break;
    default:
      if (sa.early_push) { k_fc_reduce_push<__nv_bfloat16><<<nfc, 128, 0, s>>>(ctx->last_push_fc); k_reduce_rows<__nv_bfloat16><<<nc2, 128, 0, s>>>(ctx->last_push_c2); }
      if (ctx->tail_grid) launch_k(k_step_grid<__nv_bfloat16>, dim3(STEP_GRID_CTAS), dim3(STEP_GRID_THREADS), 0, s, sa, false);
      else launch_k(k_step<__nv_bfloat16>, dim3(STEP_CTAS), dim3(STEP_THREADS), 0, s, sa, false);
      break;
  }
  CK(cudaGetLastError());
  ctx->launches += sa.early_push ? 3 : 1;
  ctx->steps_enqueued++;
  return 0;
}

// ------------------------------------------------------------------------------------ no-allocation entry point
// Some hosts of this library are not allowed to allocate device memory at the moment they
// need a result: a process that has handed its whole memory budget to another allocator (a
// framework caching pool, a server with a fixed arena carved out at start-up) and treats any
// later cudaMalloc as a bug or simply has no memory left for one, or a latency-critical loop
// where cudaMalloc's implicit device-wide synchronisation is the very thing to avoid.
// ddpstep_create() is useless to them -- it allocates a context.  For those
// callers the conv1 + BatchNorm-statistics kernel is offered on its own, with the scratch it
// needs (one row of per-channel partial sums per image, the "last CTA" ticket and the
// statistics block) living in the library image instead of in cudaMalloc'ed memory.
namespace {

constexpr int NOALLOC_MAX_BATCH = 1024;  // rows of partial sums held in the static scratch

struct NoAllocScratch {
  float part[NOALLOC_MAX_BATCH * 16 * 2];  // (sum, sum of squares) per image and channel
  BnStat st;                               // the kernel's last CTA leaves mean / invstd here
  DevState ds;                             // only ticket[0] is used; must be 0 at launch
};

// About 130 KB of global memory on every GPU that loads libddpstep.so, until the process
// exits.  In host code this name is only a stand-in: its device address comes from
// cudaGetSymbolAddress below, per device.
static __device__ __align__(256) unsigned char g_noalloc_scratch[sizeof(NoAllocScratch)];

constexpr int NOALLOC_MAX_DEVICES = 64;
std::mutex g_noalloc_mu;
void* g_noalloc_addr[NOALLOC_MAX_DEVICES] = {nullptr};

}  // namespace

extern "C" int ddpstep_conv1_bnstats_noalloc(const float* params_dev, const float* x_dev, int32_t batch, int32_t dtype,
                                             int32_t conv_bias_double_round, float bn_eps, void* y1_dev,
                                             float* stats_dev, void* stream) {
  ddpstep_ctx* ctx = nullptr;  // errors are reported through ddpstep_last_error(NULL)
  if (!params_dev || !x_dev || !y1_dev || !stats_dev) return fail(nullptr, DDPSTEP_E_ARG, "null device pointer");
  if (dtype < 0 || dtype > 2) return fail(nullptr, DDPSTEP_E_ARG, "bad dtype %d", dtype);
  // the kernel reads x with 16-byte loads and writes y1 with 8- / 16-byte stores: a misaligned sub-buffer of the
  // caller's arena would raise a sticky misaligned-address fault in the caller's whole CUDA context
  if ((reinterpret_cast<uintptr_t>(x_dev) & 15) || (reinterpret_cast<uintptr_t>(y1_dev) & 15))
    return fail(nullptr, DDPSTEP_E_ARG, "x_dev and y1_dev must be 16-byte aligned");
  if ((reinterpret_cast<uintptr_t>(params_dev) & 3) || (reinterpret_cast<uintptr_t>(stats_dev) & 3))
    return fail(nullptr, DDPSTEP_E_ARG, "params_dev and stats_dev must be 4-byte aligned");
  if (batch < 1 || batch > NOALLOC_MAX_BATCH)
    return fail(nullptr, DDPSTEP_E_ARG,
                "batch %d is outside [1, %d]: the allocation-free entry point keeps one row of partial sums per "
                "image in a fixed %zu-byte scratch array; split the batch or use a context (ddpstep_create)",
                batch, NOALLOC_MAX_BATCH, sizeof(NoAllocScratch));
  int dev = -1;
  CK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= NOALLOC_MAX_DEVICES) return fail(nullptr, DDPSTEP_E_ARG, "device ordinal %d not supported", dev);
  void* base = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (!g_noalloc_addr[dev]) {  // first call on this device: resolve the symbol with this device current
      void* p = nullptr;
      CK(cudaGetSymbolAddress(&p, g_noalloc_scratch));
      g_noalloc_addr[dev] = p;
    }
    base = g_noalloc_addr[dev];
  }
  NoAllocScratch* sc = static_cast<NoAllocScratch*>(base);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  // the array keeps whatever the previous call left: the ticket has to start at zero
  CK(cudaMemsetAsync(&sc->ds, 0, sizeof(DevState), s));
  CommDev comm;
  memset(&comm, 0, sizeof comm);
  comm.world = 1;
  Conv1Args a{BatchSrc{x_dev, nullptr, nullptr}, params_dev, y1_dev, sc->part, 0, &sc->st, &sc->ds, comm, batch, conv_bias_double_round, 0, bn_eps, nullptr, batch};
  switch (dtype) {
    case DDPSTEP_FP32: k_conv1_fwd<float><<<dim3(batch, C1_GROUPS), C1_THREADS, 0, s>>>(a); break;
    case DDPSTEP_FP16: k_conv1_fwd<__half><<<dim3(batch, C1_GROUPS), C1_THREADS, 0, s>>>(a); break;
    default: k_conv1_fwd<__nv_bfloat16><<<dim3(batch, C1_GROUPS), C1_THREADS, 0, s>>>(a); break;
  }
  CK(cudaGetLastError());
  CK(cudaMemcpyAsync(stats_dev, sc->st.lmean, 16 * sizeof(float), cudaMemcpyDeviceToDevice, s));
  CK(cudaMemcpyAsync(stats_dev + 16, sc->st.linvstd, 16 * sizeof(float), cudaMemcpyDeviceToDevice, s));
  return 0;
}

#ifdef DDPSTEP_TRACE
extern "C" int ddpstep_read_trace(unsigned long long* out64) {
  cudaDeviceSynchronize();
  return cudaMemcpyFromSymbol(out64, dstep::g_trace, 256 * 8) == cudaSuccess ? 0 : -2;
}
#endif
