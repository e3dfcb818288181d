/*
 * ddpstep.h -- C ABI of the B200-native fused DDP + AMP ConvNet training step.
 *
 * The reference (gbbin/DDP-practice) has no FFI of its own: its "operator API" is the
 * PyTorch object surface touched by ddp_main.py.  Every entry point below therefore
 * cites the reference call (file:line) whose work it performs.  The Python mirror of
 * that surface (ddp_practice_b200/shim.py) binds these symbols with ctypes; see
 * INTEGRATION.md for the binding a maintainer of the reference would add.
 *
 * Conventions: plain C types only; every function returns 0 on success or a negative
 * DDPSTEP_E_* code (ddpstep_last_error() gives the text); one host thread per context
 * (process-per-GPU, ddp_main.py:176); all device memory is owned by the library; the
 * caller supplies the CUDA stream as a void* (cudaStream_t), NULL = default stream.
 */
#ifndef DDPSTEP_H
#define DDPSTEP_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DDPSTEP_ABI_VERSION 3

#define DDPSTEP_N_PARAMS 29034   /* flat fp32 parameter / gradient bucket length   */
#define DDPSTEP_N_BNBUF 96       /* rm1[16] rv1[16] rm2[32] rv2[32]                */
#define DDPSTEP_HANDLE_BYTES 64  /* sizeof(cudaIpcMemHandle_t)                     */
#define DDPSTEP_MAX_WORLD 8

enum { DDPSTEP_FP32 = 0, DDPSTEP_FP16 = 1, DDPSTEP_BF16 = 2 };

enum {
  DDPSTEP_OK = 0,
  DDPSTEP_E_ARG = -1,      /* bad argument (batch > batch_cap, world > 8, null pointer ...) */
  DDPSTEP_E_CUDA = -2,     /* a CUDA runtime call failed                                    */
  DDPSTEP_E_STATE = -3,    /* call out of order (e.g. step before peers were imported)      */
  DDPSTEP_E_PEER = -4,     /* a peer did not answer within the timeout: the affected step and all
                              later ones applied no update; the context refuses further steps */
  DDPSTEP_E_ARCH = -5      /* device is not sm_100                                          */
};

typedef struct ddpstep_ctx ddpstep_ctx;

typedef struct ddpstep_config {
  int32_t abi_version;     /* DDPSTEP_ABI_VERSION                                            */
  int32_t rank, world;     /* dist.get_rank() / WORLD_SIZE                 ddp_main.py:65,72 */
  int32_t device;          /* CUDA device ordinal = local_rank             ddp_main.py:71    */
  int32_t dtype;           /* autocast dtype: DDPSTEP_FP16 = `autocast()`  ddp_main.py:31;
                              DDPSTEP_FP32 = origin_main.py (no autocast)                    */
  int32_t batch_cap;       /* largest per-rank batch ever passed (-b)      ddp_main.py:51    */
  int32_t use_scaler;      /* 1 = GradScaler path (ddp_main.py:91-93), 0 = origin_main.py:64 */
  int32_t sync_bn;         /* 1 = SyncBatchNorm statistics over all ranks  ddp_main.py:120   */
  int32_t use_graph;       /* 1 = replay the step as one captured CUDA graph                 */
  int32_t conv_bias_double_round; /* 1 = round conv output, then add bias and round again
                              (ATen cuDNN path); 0 = round once                              */
  float lr;                /* SGD learning rate                            ddp_main.py:125   */
  float init_scale, growth_factor, backoff_factor; /* GradScaler()  grad_scaler.py:126-128   */
  int32_t growth_interval; /*                                       grad_scaler.py:129       */
  float bn_momentum, bn_eps; /* nn.BatchNorm2d defaults             ddp_main.py:18,24        */
  /* The other knobs of torch.optim.SGD (torch/optim/sgd.py:425-465, the multi-tensor form):
   * g += weight_decay * p; buf = g on the first executed step, else buf = momentum * buf +
   * (1 - dampening) * g; g = nesterov ? g + momentum * buf : buf; p -= lr * g.  All zero
   * (the reference, ddp_main.py:125) leaves the plain update.  A step skipped by the scaler
   * touches neither p nor buf.  SURVEY.md section 8(f) N3.                                  */
  float momentum, weight_decay, dampening;
  int32_t nesterov;        /* needs momentum > 0 and dampening == 0, as in torch             */
} ddpstep_config;

/* Device pointers into the context's state, for zero-copy views (torch tensors that alias
 * them back `model.parameters()`, `p.grad`, BN buffers, `scaler._scale`).                    */
typedef struct ddpstep_ptrs {
  float* params;           /* [29034] fp32 master weights, definition order  ddp_main.py:16-28 */
  float* grads;            /* [29034] tap T-c: all-reduced, unscaled grads   ddp_main.py:92    */
  float* grads_scaled;     /* [29034] tap T-b: all-reduced, still scaled     ddp_main.py:91    */
  float* bn_buffers;       /* [96] running_mean/var of both BN layers                          */
  int64_t* num_batches_tracked; /* [2]                                                         */
  float* scale;            /* [1] GradScaler._scale                          grad_scaler.py:176 */
  int32_t* growth_tracker; /* [1] GradScaler._growth_tracker                                   */
  float* found_inf;        /* [1] of the last step (0.0 / 1.0)               grad_scaler.py:371 */
  float* loss;             /* [1] tap T-a: this rank's loss of the last step ddp_main.py:89    */
  void* logits;            /* [batch_cap,10] in `dtype`: outputs of the last forward           */
  float* counts;           /* [2] correct, size accumulators of test()       ddp_main.py:99-107 */
  float* input_stage;      /* [batch_cap,784] fp32 staging buffer for images                   */
  int64_t* label_stage;    /* [batch_cap] staging buffer for labels                            */
  float* momentum_buffer;  /* [29034] SGD momentum buffers (optimizer.state[p]['momentum_buffer']) */
  int32_t* momentum_ready; /* [1] 1 once a step has initialised them (torch creates them lazily) */
} ddpstep_ptrs;

/* Fill a config with the reference's defaults (fp16 autocast, GradScaler(), lr 1e-4). */
void ddpstep_default_config(ddpstep_config* cfg);

/* ConvNet().cuda() + DDP(...) + SGD + GradScaler() construction: ddp_main.py:117-126.
 * Allocates all device state for one rank on cfg->device.  Parameters start at zero:
 * load them with ddpstep_load_state.                                                        */
int ddpstep_create(const ddpstep_config* cfg, ddpstep_ctx** out);
/* dist.destroy_process_group() + object teardown: ddp_main.py:170.                          */
int ddpstep_destroy(ddpstep_ctx* ctx);
const char* ddpstep_last_error(const ddpstep_ctx* ctx);  /* ctx may be NULL: create() errors */

/* Peer bootstrap, replaces ProcessGroupNCCL communicator setup (init_process_group,
 * ddp_main.py:73).  export: 64-byte IPC handle of this rank's symmetric region; the caller
 * all-gathers the handles over its own control plane (the c10d store) and passes the
 * world*64 bytes, rank-major, to import.  world == 1 needs neither.                         */
int ddpstep_export_handle(ddpstep_ctx* ctx, void* handle64);
int ddpstep_import_peers(ddpstep_ctx* ctx, const void* handles, int32_t world);

/* DDP's construction-time broadcast of rank 0's parameters and buffers
 * (distributed.py:881-890) ends with every rank calling this with the same host arrays.
 * nbt may be NULL (zeros).                                                                  */
int ddpstep_load_state(ddpstep_ctx* ctx, const float* params29034, const float* bn_buffers96,
                       const int64_t* nbt2);
/* model.state_dict() / scaler.state_dict(): ddp_main.py:165-169.  Any pointer may be NULL.  */
int ddpstep_read_state(ddpstep_ctx* ctx, float* params29034, float* bn_buffers96, int64_t* nbt2,
                       float* scale, int32_t* growth_tracker);
int ddpstep_set_scaler(ddpstep_ctx* ctx, float scale, int32_t growth_tracker);
int ddpstep_state_ptrs(ddpstep_ctx* ctx, ddpstep_ptrs* out);

/* One iteration of train(): ddp_main.py:86-93 (origin_main.py:60-66 when use_scaler == 0):
 * forward, CrossEntropyLoss, scale(loss).backward() with the DDP all-reduce, scaler.step
 * (unscale, inf check, SGD) and scaler.update, as one stream-ordered launch sequence /
 * graph replay.  x: [batch,1,28,28] fp32, y: [batch] int64, both DEVICE pointers.          */
int ddpstep_train_step(ddpstep_ctx* ctx, const float* x_dev, const int64_t* y_dev, int32_t batch,
                       void* stream);
/* K consecutive iterations of train() (the `for images, labels in train_dloader` loop of
 * ddp_main.py:85-93) as ONE graph launch.  x_dev [n_steps*batch,1,28,28] fp32 and y_dev
 * [n_steps*batch] int64 hold n_steps consecutive batches on the device.  Identical arithmetic
 * and state to n_steps calls of ddpstep_train_step; inside the graph the steps are chained by
 * programmatic edges, so the gap between two graph launches and the host's launch jitter (which
 * at world > 1 the first exchange of every step has to absorb) are paid once per call.          */
int ddpstep_train_steps(ddpstep_ctx* ctx, const float* x_dev, const int64_t* y_dev, int32_t n_steps,
                        int32_t batch, void* stream);
/* One iteration inside `with model.no_sync():` -- gradient accumulation, SURVEY.md 8(f) N3
 * (torch/nn/parallel/distributed.py:1441-1470; the reference never uses it): forward, loss and
 * backward as in ddpstep_train_step (BN statistics, running buffers and SyncBN exchanges
 * included), but the local gradient is added to an accumulation buffer -- what `.grad` does
 * when zero_grad() is not called -- and nothing is all-reduced, unscaled or applied.  The next
 * ddpstep_train_step* call reduces and applies accumulated + own gradient.                   */
int ddpstep_train_step_accumulate(ddpstep_ctx* ctx, const float* x_dev, const int64_t* y_dev,
                                  int32_t batch, void* stream);
/* Same, from HOST buffers (the `.cuda()` copies of ddp_main.py:86-87 included).  Pinned
 * buffers (cudaHostAlloc / pin_memory()) are copied by two nodes of the step graph itself,
 * pageable ones by ordinary stream-ordered copies.  If loss_out is not NULL the call returns
 * once the step has posted its loss (the tail kernel writes it into a pinned mailbox the call
 * polls: no device-to-host copy, no stream synchronisation); with loss_out == NULL the call
 * is asynchronous and, as with cudaMemcpyAsync, the host buffers must stay untouched until
 * the stream has passed the step.                                                           */
int ddpstep_train_step_host(ddpstep_ctx* ctx, const float* x_host, const int64_t* y_host,
                            int32_t batch, float* loss_out, void* stream);

/* The input pipeline in front of the step: instead of torchvision MNIST -> ToTensor ->
 * DataLoader workers -> pinned batch -> `.cuda()` (ddp_main.py:127-141, 86-87), the dataset stays in HBM as
 * the uint8 bytes MNIST consists of, images_u8_dev [N,784] and labels_dev [N], and a step receives the
 * DistributedSampler's indices for this batch, idx_dev [batch] (torch/utils/data/distributed.py:107-141).
 * The gather and ToTensor's `uint8 -> fp32, / 255` (torchvision/transforms/functional.py:157) happen inside
 * conv1's load.  Same arithmetic, bit for bit, as ddpstep_train_step on images_u8[idx].float() / 255.        */
int ddpstep_train_step_indexed(ddpstep_ctx* ctx, const uint8_t* images_u8_dev, const int64_t* labels_dev,
                               const int64_t* idx_dev, int32_t batch, void* stream);
/* ... and an epoch segment of it in one launch: idx_dev holds n_steps*batch sampler indices
 * (what `train_dloader.sampler` yields after set_epoch, ddp_main.py:160), n_steps full batches.  */
int ddpstep_train_steps_indexed(ddpstep_ctx* ctx, const uint8_t* images_u8_dev, const int64_t* labels_dev,
                                const int64_t* idx_dev, int32_t n_steps, int32_t batch, void* stream);
int ddpstep_eval_batch_indexed(ddpstep_ctx* ctx, const uint8_t* images_u8_dev, const int64_t* labels_dev,
                               const int64_t* idx_dev, int32_t batch, void* stream);

/* One iteration of test(): ddp_main.py:101-107.  Eval-mode forward (running statistics),
 * logits left in ptrs.logits; when y_dev != NULL also `size += batch` and
 * `correct += (argmax == y).sum()` into ptrs.counts.                                        */
int ddpstep_eval_batch(ddpstep_ctx* ctx, const float* x_dev, const int64_t* y_dev, int32_t batch,
                       void* stream);
int ddpstep_reset_counts(ddpstep_ctx* ctx, void* stream);        /* ddp_main.py:99-100       */
/* dist.reduce(size, 0, SUM); dist.reduce(correct, 0, SUM): ddp_main.py:108-109.  On rank 0
 * ptrs.counts then holds the sums over all ranks.                                           */
int ddpstep_reduce_counts(ddpstep_ctx* ctx, void* stream);
int ddpstep_read_counts(ddpstep_ctx* ctx, float* correct, float* size, void* stream);

/* Test / profiling taps. which: 0 y1, 1 p1, 2 y2, 3 p2 (activations in `dtype`),
 * 4 dlogits, 5 d(conv2 out), 6 d(pool1 out).  Copies min(bytes, size) to host.             */
int ddpstep_read_activation(ddpstep_ctx* ctx, int32_t which, void* host, int64_t bytes);
/* Profiling aid for bench.py: run ONE training step (state advances like any step) with a
 * CUDA event after each of its kernels (ddpstep_kernel_name lists them; n_out >= 16 is always
 * enough) and return their device durations in ms.
 * in_graph = 1 replays the step as a graph with event-record nodes between the kernel nodes
 * (what the hot path does), 0 launches eagerly.  A NEGATIVE entry marks a kernel that does
 * not run at this batch size (its work rides in another kernel: the filter-tile refresh in
 * conv1, BN2's backward reduction in Linear+CE); its magnitude is the bare interval between
 * two consecutive event records, i.e. the instrumentation overhead contained in every other
 * entry.  No reference counterpart: the reference has no per-kernel instrumentation
 * (SURVEY.md section 5).                                                                    */
int ddpstep_profile_step(ddpstep_ctx* ctx, const float* x_dev, const int64_t* y_dev, int32_t batch,
                         int32_t in_graph, float* ms_out, int32_t n_out, void* stream);
const char* ddpstep_kernel_name(int32_t i);   /* names of those kernels, "" past the end */
/* conv1 (5x5, pad 2, 1->16 channels) plus the per-channel BatchNorm batch statistics of its
 * output, WITHOUT a context and without allocating anything.  No counterpart in the
 * reference: it exists for hosts that cannot call cudaMalloc when they need the result -- a
 * process whose device memory all belongs to another allocator (a caching pool, a fixed arena
 * taken at start-up), or a loop that must not hit cudaMalloc's device-wide synchronisation;
 * such a caller cannot use ddpstep_create, which allocates a context.  Everything
 * is caller-owned device memory: params_dev = the first 416 floats of the flat parameter
 * vector (conv1 weight[400], bias[16]); x_dev = [batch,1,28,28] fp32; y1_dev = [batch,16,28,28]
 * in `dtype`; stats_dev = 32 floats, mean[16] then 1/sqrt(biased var + bn_eps)[16].
 * Alignment: x_dev and y1_dev 16 bytes (the kernel uses 16-byte loads and 8/16-byte stores),
 * params_dev and stats_dev 4 bytes; anything else is refused with DDPSTEP_E_ARG rather than
 * faulting inside the caller's CUDA context.
 * Limit: batch <= 1024 (DDPSTEP_E_ARG beyond it, text via ddpstep_last_error(NULL)): the
 * kernel's scratch -- a row of partial sums per image -- is a fixed array of about 130 KB
 * inside the library, resident on every GPU that loads it until the process exits.
 * The target device must be current.  The first call on each device looks that array's
 * address up; make that first call before starting any stream capture.  All calls share the
 * one array per device and it keeps the previous call's contents (the call re-zeroes the part
 * that must start at zero), so two calls whose work overlaps on the GPU -- different streams,
 * or different threads -- race; order them on one stream or with events.  Asynchronous. */
int ddpstep_conv1_bnstats_noalloc(const float* params_dev, const float* x_dev, int32_t batch, int32_t dtype,
                                  int32_t conv_bias_double_round, float bn_eps, void* y1_dev, float* stats_dev,
                                  void* stream);
/* `optimizer.param_groups[0]` at run time -- an LR scheduler's step, or the group edited by hand
 * (torch/optim/sgd.py:425-465 reads lr / momentum / weight_decay / dampening / nesterov from the
 * group on every optimizer.step(); ddp_main.py:125 sets them once).  The values live in device
 * memory, so neither the context nor its captured graphs are rebuilt.  Stream-ordered: steps
 * enqueued on `stream` after this call use the new values.                                    */
int ddpstep_set_sgd(ddpstep_ctx* ctx, float lr, float momentum, float weight_decay, float dampening,
                    int32_t nesterov, void* stream);
/* The same for ONE of several `optimizer.param_groups` (torch/optim/sgd.py:425-465 runs once per
 * group): tensor_mask selects the parameter tensors of the group, bit i = tensor i in definition
 * order (0 conv1.weight, 1 conv1.bias, 2 bn1.weight, 3 bn1.bias, 4 conv2.weight, 5 conv2.bias,
 * 6 bn2.weight, 7 bn2.bias, 8 fc.weight, 9 fc.bias).  ddpstep_set_sgd == mask 0x3FF.  SURVEY 8f N3. */
int ddpstep_set_sgd_group(ddpstep_ctx* ctx, uint32_t tensor_mask, float lr, float momentum,
                          float weight_decay, float dampening, int32_t nesterov, void* stream);
/* How long a kernel polls for a peer's words before giving up (default 20 s; environment
 * DDPSTEP_PEER_TIMEOUT_MS at create time).  The reference's NCCL blocks until its watchdog
 * fires; here the step whose wait expired applies no update and this and
 * every later step entry point returns DDPSTEP_E_PEER.                                        */
int ddpstep_set_peer_timeout_ms(ddpstep_ctx* ctx, double ms);
/* The Reducer all-reduce of the one 116,136-byte bucket with everything fused behind it
 * (ncclAllReduce of reducer.hpp:524 behind ddp_main.py:121-123, then grad_scaler.py:294-298,371
 * and sgd.py:473) launched ALONE: the tail kernel re-reduces the partial gradients the last
 * training step left behind (state advances as for a step).  Exists so that the exchange can be
 * timed in isolation against ncclAllReduce (tools/allreduce_bench.py, bench.py `allreduce`);
 * collective: every rank calls it the same number of times.                                   */
int ddpstep_allreduce_only(ddpstep_ctx* ctx, void* stream);
/* The work of a DDP communication hook for one GradBucket of ANY model -- the plugin API the
 * reference's stack already has: `DistributedDataParallel.register_comm_hook(state, hook)`,
 * hook(state, bucket) -> Future[Tensor] (torch/nn/parallel/distributed.py:1987-2067; C++
 * CommHookInterface::runHook, comm.hpp:104-118; SURVEY.md 8b alternative B, 8f N3).
 * bucket <- sum over ranks of bucket * (1/W) if `average` (what DDP's default hook computes,
 * default_comm_hooks.hpp:36-51), plain sum otherwise; in place; identical bits on every rank
 * (fixed rank-order sum).  Any length: chunks of 29,056 elements go through the step's own
 * double-buffered inbox.  Collective, stream-ordered.  ddp_practice_b200/comm_hook.py wraps it. */
int ddpstep_allreduce_bucket(ddpstep_ctx* ctx, float* bucket_dev, int64_t n, int32_t average, void* stream);
/* Number of kernel launches (graph kernel nodes count one each) issued so far.             */
int64_t ddpstep_launch_count(const ddpstep_ctx* ctx);
/* Device-side error word (peer timeout), 0 = healthy.  Synchronises the device.  The step
 * entry points report the same condition without synchronising, one or two steps late.      */
int ddpstep_device_status(ddpstep_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif /* DDPSTEP_H */
