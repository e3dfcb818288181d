#!/usr/bin/env python
"""Benchmark of the hot path: one ConvNet training step under DDP + AMP (BASELINE.json).

  python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
  python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU path (host cores)
  torchrun --nproc-per-node N ... bench.py --gpus N ...    # N > 1, one rank per GPU

Workload (config.workload): BASELINE.json configs[1] -- "ddp_main.py ConvNet with AMP autocast
+ GradScaler, batch 32/GPU" (fp16, SyncBN + DDP all-reduce when N > 1).  A step is one pass of
ddp_main.py:86-93 over one batch of synthetic MNIST-shaped input.  Prints ONE JSON line:

  value / ms_per_step   K steps, inputs resident in HBM, through ddpstep_train_steps (min(K, 50) steps per graph launch)
  per_step_launch       the same K steps as one ddpstep_train_step call (one graph launch) each; p50_step_us likewise
  e2e                   pinned HOST batch in, loss out, every step, through ddpstep_train_step_host
  roofline              dominant kernel (event nodes inside the step graph), latency floor, ncu dram traffic
  parity_checks         GradScaler state vs the reference's data-independent trajectory; replicas bit-identical (N > 1)
  allreduce (N > 1)     the 116,136-byte bucket: bare exchange, fused tail, eager NCCL
  large_batch           BASELINE.json configs[3]: bf16, batch 4096 per GPU, 20 steps
  input_pipeline        the step fed from a GPU-resident uint8 dataset by DistributedSampler indices
  cpu_baseline (N = 1)  origin_main.py's fp32 step on the host cores (also `--impl reference`)
Timing scheme: DESIGN.md section 7."""
import argparse
import json
import os
import statistics
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "ConvNet train imgs/sec (max over ranks)"
UNIT = "img/s"
N_TRAIN = 60000  # MNIST-sized resident dataset: 188 MB fp32 > the 126 MB L2

# algorithmic work per image of each kernel: (flops, activation ELEMENTS moved, fp32-input bytes)
# (DESIGN.md "Kernels" states the same table)
KERNEL_WORK = {
    "conv1_fwd": (627200, 12544, 784 * 4), "bn1_relu_pool": (0, 12544 + 3136, 0),
    "conv2_fwd": (5017600, 3136 + 6272, 0), "bn2_relu_pool": (0, 6272 + 1568, 0),
    "fc_ce": (62720, 1568 + 1568, 0), "fc_wgrad": (31360, 1568, 0),
    "bn2_bwd_reduce": (0, 6272 + 1568, 0), "bn2_bwd_apply": (0, 6272 * 2 + 1568, 0),
    "conv2_wgrad": (5017600, 3136 + 6272, 0), "conv2_dgrad": (5017600, 6272 + 3136, 0),
    "bn1_bwd_reduce": (0, 12544 + 3136, 0), "conv1_wgrad": (627200, 12544 + 3136, 784 * 4),
    "step": (0, 0, 0), "conv2_wgrad_reduce": (0, 0, 0), "w2_tiles_refresh": (0, 0, 0),
}
STEP_PARAM_BYTES = 4 * 116136  # params read+write, grads write x2, per step (not per image)
STEP_BYTES_PER_IMG_AMP = 247744  # SURVEY.md 8(d), materialise-once model
STEP_FLOPS_PER_IMG = 16401280


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d["hbm_gbs"], d["bf16_tflops"], "measured (MEASURED_PEAKS.json, burst)"
    return 6650.0, 1590.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """Samples SM clock and throttle reasons while the timed region runs."""

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._t = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        nv = self.nv
        names = {nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
                 nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap"}
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.02)

    def __enter__(self):
        if self.nv:
            self._t = threading.Thread(target=self._run, daemon=True)
            self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._t:
            self._t.join()

    def summary(self):
        return {"sm_mhz": statistics.median(self.samples) if self.samples else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


def cpu_reference_run(batch, steps, warmup, budget_s, threads=0):
    """Times oracle/reference_step.py (the reference's CPU path) on this host's cores."""
    import torch
    from oracle.convnet_oracle import synthetic_mnist
    from oracle.reference_step import CpuReference
    ref = CpuReference(threads)
    x, y = synthetic_mnist(batch * min(64, max(4, 8192 // batch)))
    times, loss = ref.timed(x, y, batch, steps, warmup, budget_s)
    sec = sum(times) / len(times)
    return {"value": batch / sec, "unit": UNIT, "cores": ref.threads, "kind": "port",
            "sample": f"{len(times)} steps of origin_main.py's fp32 step, batch {batch}, after {warmup} warm-up; "
                      f"torch {torch.__version__} CPU kernels, {ref.threads} threads of {os.cpu_count()} host cpus",
            "ms_per_step": sec * 1e3, "p50_ms": statistics.median(times) * 1e3}


def run_reference(args):
    """The reference arm.  N ranks of DDP + SyncBN on batch B each compute exactly the global-batch step
    (SURVEY.md finding 6), so at --gpus N the reference's CPU form of the same work is ONE origin_main.py step on the
    global batch N*B; it runs on rank 0 with every host thread it can use (torchrun pins OMP_NUM_THREADS=1, which
    is undone here), the other ranks exit."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    gb = args.batch * args.gpus
    b = cpu_reference_run(gb, args.steps, args.warmup, budget_s=240.0, threads=os.cpu_count() or 1)
    line = {"impl": "reference", "metric": METRIC, "value": b["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": b["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args, cpu=True),
            "cpu_baseline": {k: b[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": b["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def workload_config(args, cpu=False):
    return {"workload": f"BASELINE.json configs[1]: ddp_main.py ConvNet step, AMP {args.dtype} autocast + GradScaler, "
                        f"batch {args.batch}/GPU, SyncBN + one-bucket DDP all-reduce when n_gpus > 1"
                        + (" -- reference arm: the reference's CPU-runnable form of the same step (configs[0], "
                           f"origin_main.py fp32 SGD on the host cores) on the GLOBAL batch of {args.batch * args.gpus}, "
                           "which is what N ranks of DDP + SyncBN compute together" if cpu else ""),
            "batch_per_gpu": args.batch, "global_batch": args.batch * args.gpus,
            "l2": f"inputs larger than L2: {N_TRAIN} resident images ({N_TRAIN * 784 * 4 / 1e6:.0f} MB fp32), "
                  "each batch read once per pass; the 2 MB of step activations are L2-resident by construction",
            "parallelism": f"dp{args.gpus}"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=200)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--dtype", default="fp16", choices=["fp16", "bf16", "fp32"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-large-batch", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    from ddp_practice_b200 import Engine  # fails loudly without the CUDA library / a GPU

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torchrun --nproc-per-node {args.gpus}")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    B, K, W = args.batch, args.steps, args.warmup

    # ---- synthetic dataset, resident in HBM, sharded like DistributedSampler (rank::world)
    g = torch.Generator().manual_seed(3407)
    n_local = (N_TRAIN // world // B) * B
    x_all = (torch.randint(0, 256, (n_local, 1, 28, 28), dtype=torch.uint8, generator=g).float() / 255)
    y_all = torch.randint(0, 10, (n_local,), generator=g)
    if world > 1:  # different data per rank
        x_all = x_all.roll(rank * 997, 0)
        y_all = y_all.roll(rank * 997, 0)
    xd, yd = x_all.cuda(), y_all.cuda()
    nb = n_local // B

    from oracle.reference_step import build_model  # random-init weights of the reference architecture
    from oracle.convnet_oracle import OracleState
    st = OracleState.from_state_dict(build_model().state_dict())
    eng = Engine(dtype=args.dtype, batch_cap=B, rank=rank, world=world, device=local, use_graph=not args.no_graph,
                 use_scaler=args.dtype != "fp32")
    eng.connect()
    bn = torch.cat([st.running_mean[0], st.running_var[0], st.running_mean[1], st.running_var[1]])
    eng.load_state(st.flat_params().numpy(), bn.numpy())

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v):
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item()

    def batch(i):
        j = i % nb
        return xd[j * B:(j + 1) * B], yd[j * B:(j + 1) * B]

    # ---- device-resident timing: W warm-up, host barrier + synchronize, then ALIGN more untimed steps enqueued
    # back-to-back with exactly K timed ones, the first event recorded on the stream between the two groups.
    # Why the alignment steps: after a host barrier every rank re-starts its GPU from the host at a slightly different
    # moment, and the first cross-GPU exchange of the first step absorbs that skew (57 us in one kernel at
    # N=8).  Every step contains five exchanges, so after one step the GPUs run in lock-step; the skew therefore lands
    # in the untimed alignment steps and the timed region measures steady-state steps only, also at --steps 20.
    ALIGN = 3
    steps_done = 0

    def run_steps(n, first):
        for i in range(n):
            eng.train_step(*batch(first + i))

    run_steps(W, 0)
    steps_done += W
    barrier()
    launches0 = eng.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clk:
        run_steps(ALIGN, W)
        launches_align = eng.launch_count() - launches0
        ev0.record()
        run_steps(K, W + ALIGN)
        ev1.record()
        barrier()
    steps_done += ALIGN + K
    ms_single = max_over_ranks(ev0.elapsed_time(ev1))
    launches_single = eng.launch_count() - launches0 - launches_align
    eng.check_device()

    # ---- the same K steps through ddpstep_train_steps: CHUNK consecutive steps per graph launch (chained inside the
    # graph by programmatic edges).  Same kernels, same arithmetic, bit-identical state (tests/test_gpu_parity.py::
    # test_multi_step_graph_equals_single_step_launches); what goes away is the gap between two graph launches and the
    # host's per-launch jitter.  This is the device-resident headline (`value`); the per-launch figure stays beside it.
    CHUNK = min(K, 50)

    def run_chunks(n, first):
        done = 0
        while done < n:
            k = min(CHUNK, n - done)
            j = (first + done) % nb
            if j + k > nb:
                j = 0
            eng.train_steps(xd[j * B:(j + k) * B], yd[j * B:(j + k) * B], B)
            done += k

    run_chunks(min(W, CHUNK), 0)  # warm-up; also captures every graph size used below, outside the timed region
    run_chunks(CHUNK, 0)
    run_chunks(ALIGN, 0)
    if K % CHUNK:
        run_chunks(K % CHUNK, 0)
    steps_done += min(W, CHUNK) + CHUNK + ALIGN + (K % CHUNK)
    barrier()
    launches1 = eng.launch_count()
    m0, m1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clk2:
        run_chunks(ALIGN, W)
        launches_align = eng.launch_count() - launches1
        m0.record()
        run_chunks(K, W + ALIGN)
        m1.record()
        barrier()
    steps_done += ALIGN + K
    ms_total = max_over_ranks(m0.elapsed_time(m1))
    gpu_launches = eng.launch_count() - launches1 - launches_align
    eng.check_device()

    # ---- cheap parity evidence on every run (also under torchrun): (1) GradScaler's state after `steps_done` steps
    # equals the data-independent trajectory of the reference (SURVEY.md finding 2/7: fp16 overflows at step 0 and
    # once after every growth to 65536; bf16 never) -- bit-exact; (2) at N > 1 every replica holds bit-identical
    # parameters, BN buffers and scaler state (DDP's invariant).
    def expected_scaler(n, dtype):
        scale, tr = 65536.0, 0
        for _ in range(n):
            if dtype == "fp16" and scale >= 65536.0:
                scale, tr = scale * 0.5, 0
            else:
                tr += 1
                if tr == 2000:
                    scale, tr = scale * 2.0, 0
        return scale, tr

    checks = {}
    if args.dtype != "fp32":
        got = (eng.scale.item(), eng.growth_tracker.item())
        checks["scaler_state_after_steps"] = {"steps": steps_done, "got": list(got),
                                              "expected": list(expected_scaler(steps_done, args.dtype)),
                                              "ok": got == expected_scaler(steps_done, args.dtype)}
    if world > 1:
        mine = torch.cat([eng.params, eng.bn_buffers, eng.scale, eng.growth_tracker.float()]).clone().view(torch.int32)
        ref0 = mine.clone()
        dist.broadcast(ref0, 0)
        same = torch.tensor([int(torch.equal(mine, ref0))], device="cuda")
        dist.all_reduce(same, op=dist.ReduceOp.MIN)
        checks["replicas_bit_identical"] = bool(same.item())
        if not checks["replicas_bit_identical"]:
            raise SystemExit("replicas diverged: parameters / BN buffers / scaler state differ between ranks")
    checks["params_finite"] = bool(torch.isfinite(eng.params).all().item())

    # ---- p50 step latency: same loop with an event per step (after the same alignment)
    n_lat = min(K, 1000)
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(n_lat + 1)]
    barrier()
    run_steps(ALIGN, 0)
    evs[0].record()
    for i in range(n_lat):
        eng.train_step(*batch(ALIGN + i))
        evs[i + 1].record()
    barrier()
    lat = [evs[i].elapsed_time(evs[i + 1]) * 1e3 for i in range(n_lat)]
    p50_us = max_over_ranks(statistics.median(lat))
    # what the alignment steps absorb: a step launched right after a host barrier against the median step
    evs2 = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    barrier()
    evs2[0].record()
    eng.train_step(*batch(0))
    evs2[1].record()
    barrier()
    skew_us = max_over_ranks(evs2[0].elapsed_time(evs2[1]) * 1e3) - p50_us

    # ---- end to end through the public host API: pinned host batch in, loss out, every step
    nh = min(64, nb)
    xh, yh = x_all[: B * nh].pin_memory(), y_all[: B * nh].pin_memory()
    k_e2e = min(K, 1000)
    host_batches = [(xh[j * B:(j + 1) * B], yh[j * B:(j + 1) * B]) for j in range(nh)]  # what a pinned DataLoader yields
    for i in range(10):
        eng.train_step_host(*host_batches[i % nh])
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for i in range(ALIGN):
        eng.train_step_host(*host_batches[i % nh], want_loss=True)
    e0.record()
    for i in range(k_e2e):
        eng.train_step_host(*host_batches[(ALIGN + i) % nh], want_loss=True)
    e1.record()
    barrier()
    e2e_ms = max_over_ranks(e0.elapsed_time(e1))

    # ---- the input pipeline fused in -- resident uint8 dataset + DistributedSampler indices
    from torch.utils.data.distributed import DistributedSampler
    u8 = (x_all * 255).round().to(torch.uint8).cuda()
    sampler = DistributedSampler(range(n_local), num_replicas=1, rank=0)
    sampler.set_epoch(0)
    order = torch.tensor(list(iter(sampler)), dtype=torch.int64).cuda()
    idx_batches = [order[j * B:(j + 1) * B].contiguous() for j in range(nb)]
    for i in range(W):
        eng.train_step_indexed(u8, yd, idx_batches[i % nb])
    barrier()
    p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for i in range(ALIGN):
        eng.train_step_indexed(u8, yd, idx_batches[i % nb])
    p0.record()
    for i in range(K):
        eng.train_step_indexed(u8, yd, idx_batches[(W + i) % nb])
    p1.record()
    barrier()
    pipe_ms = max_over_ranks(p0.elapsed_time(p1))

    # ---- the DDP all-reduce in isolation (north_star: "allreduce latency and bus GB/s against NVLink peak"): the tail
    # kernel alone (exchange + unscale + inf check + SGD + scale update in one launch) against ncclAllReduce on the same
    # 116,136-byte bucket, both back-to-back on one stream, device-timed, max over ranks.  tools/allreduce_bench.py is
    # the long form (adds symmetric-memory one-shot and the SyncBN-sized exchanges).
    allreduce = None
    if world > 1:
        n_ar, per = 40, 25

        def timed(fn, per_call):
            for _ in range(5):
                fn()
            barrier()
            a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            for _ in range(ALIGN):
                fn()
            a0.record()
            for _ in range(n_ar):
                fn()
            a1.record()
            barrier()
            return max_over_ranks(a0.elapsed_time(a1)) / n_ar / per_call * 1e3

        def graph_of(fn):  # `per` calls of OUR kernels in one CUDA graph: device-side cost, no host launch overhead
            g, sg = torch.cuda.CUDAGraph(), torch.cuda.Stream()
            with torch.cuda.stream(sg):
                fn()
                torch.cuda.synchronize()
                with torch.cuda.graph(g, stream=sg):
                    for _ in range(per):
                        fn()
            return g

        bucket = torch.zeros(29034, device="cuda")
        mine = torch.zeros(29034, device="cuda")
        barrier()
        g_bare, g_tail = graph_of(lambda: eng.allreduce_bucket(mine)), graph_of(eng.allreduce_only)
        bare_us = timed(g_bare.replay, per)
        tail_us = timed(g_tail.replay, per)
        nccl_us = timed(lambda: dist.all_reduce(bucket), 1)
        del g_bare, g_tail
        S = 116136
        bus = lambda us: S / (us * 1e-6) / 1e9 * 2 * (world - 1) / world
        allreduce = {"bytes": S, "latency_us": bare_us, "busbw_gbs": bus(bare_us), "algbw_gbs": S / (bare_us * 1e-6) / 1e9,
                     "fused_tail_us": tail_us, "nccl_latency_us": nccl_us, "nccl_busbw_gbs": bus(nccl_us),
                     "nvlink_peak_gbs_per_dir": 900.0,
                     "what": f"latency_us: ddpstep_allreduce_bucket on the 116,136-byte bucket (the bare one-shot exchange), {per} calls per "
                             f"CUDA-graph replay x {n_ar} replays; fused_tail_us: ddpstep_allreduce_only (exchange + unscale + non-finite "
                             "check + SGD + scale update, what the step runs); nccl_latency_us: dist.all_reduce of the same bucket "
                             "issued eagerly, as the reference's Reducer issues it (reducer.hpp:524).  Device-timed, max over ranks.  "
                             "The message is latency-bound: 116 KB is 0.13 us of NVLink wire time.  tools/allreduce_bench.py adds NCCL "
                             "inside a CUDA graph, torch's symmetric-memory one-shot and the NVLS multimem kernel "
                             "(profiles/r02_allreduce_w*.json)"}
        eng.check_device()

    # ---- per-kernel device times (graph with event nodes), for the roofline of the dominant kernel
    prof = {}
    n_prof = 20
    for i in range(n_prof):
        for k, v in eng.profile_step(*batch(i), in_graph=True).items():
            prof[k] = prof.get(k, 0.0) + v / n_prof
    barrier()
    hbm_peak, tf_peak, peak_src = peaks()
    es = 4 if args.dtype == "fp32" else 2
    roof = roofline_of(prof, B, es, hbm_peak, tf_peak, peak_src, n_prof)
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "r02_traffic.json")
    if os.path.exists(tpath) and B == 32 and args.dtype == "fp16":  # the capture was taken on this configuration
        traffic = json.load(open(tpath)).get(roof["kernel"])
    roof["traffic"] = traffic
    roof["step_hbm_frac"] = (STEP_BYTES_PER_IMG_AMP * (es / 2) * B + STEP_PARAM_BYTES) / (ms_total / K * 1e-3) / 1e9 / hbm_peak
    roof["note"] = "batch 32 is latency-bound (SURVEY.md 8d): 7.9 MB and 0.52 GFLOP per step"
    # SURVEY 8(d)'s stated floor for the latency-bound case: kernel nodes on the critical path of the step graph x the
    # measured per-node gap + one graph launch + the cross-GPU exchanges x one NVLink store-and-poll round trip
    # (profiles/r01_microbench_launch_floor.txt; 0.7 us per PDL node, 2.05 us per graph launch, ~2.2 us per exchange)
    crit_nodes = 10 if B <= 64 else 11
    roof["latency_floor_us"] = {"critical_path_nodes": crit_nodes, "per_node_us": 0.7, "graph_launch_us": 2.05,
                                "exchanges": 5 if world > 1 else 0, "per_exchange_us": 2.2,
                                "floor_us": round(crit_nodes * 0.7 + 2.05 + (5 * 2.2 if world > 1 else 0.0), 2),
                                "what": "launch/exchange floor only; the measured step also pays 1-2 L2 round trips per stage"}
    eng.close()

    # ---- BASELINE.json configs[3]: bf16 autocast, batch 4096 per GPU, at this N (the regime where the convolutions are
    # dense contractions).  Same timing scheme; a short run (about 3 s) so the headline stays the configs[1] number.
    large = None
    if not args.no_large_batch and args.batch == 32:
        large = large_batch_run(torch, dist, Engine, st, rank, world, local, barrier, max_over_ranks, hbm_peak, tf_peak, peak_src, ALIGN)

    line = {"metric": METRIC, "value": world * B * K / (ms_total * 1e-3), "unit": UNIT, "n_gpus": world, "steps": K,
            "warmup": W, "ms_per_step": ms_total / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": {"fp16": "f16", "bf16": "bf16", "fp32": "f32"}[args.dtype] + " storage, f32 accumulate",
            "data": "synthetic", "config": dict(workload_config(args), align_steps=ALIGN,
                                                steps_per_graph_launch=CHUNK,
                                                timing="W warm-up steps, host barrier + synchronize, "
                                                       f"{ALIGN} untimed alignment steps, then K steps between two CUDA events "
                                                       "enqueued back-to-back, host barrier + synchronize; max over ranks; "
                                                       f"`value`: ddpstep_train_steps, {CHUNK} steps per graph launch; "
                                                       "`per_step_launch`: one graph launch per step; p50_step_us: per-step launches"),
            "p50_step_us": p50_us, "skew_us": skew_us,
            "per_step_launch": {"value": world * B * K / (ms_single * 1e-3), "unit": UNIT, "ms_per_step": ms_single / K,
                                "gpu_launches": launches_single, "clocks": clk.summary(),
                                "what": "the same K steps as one ddpstep_train_step call (one graph launch) per step"},
            "clocks": clk2.summary(),
            "e2e": {"value": world * B * k_e2e / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": B * 784 * 4 + B * 8,
                    "d2h_bytes_per_step": 4, "steps": k_e2e, "ms_per_step": e2e_ms / k_e2e,
                    "api": "ddpstep_train_step_host: pinned host batch copied in, loss returned through a pinned mailbox, every step"},
            "gpu_launches": gpu_launches, "roofline": roof, "parity_checks": checks,
            "input_pipeline": {"value": world * B * K / (pipe_ms * 1e-3), "unit": UNIT, "ms_per_step": pipe_ms / K,
                               "what": "same step fed from a GPU-resident uint8 dataset by DistributedSampler indices; "
                                       "gather and ToTensor's /255 fused into conv1's load (ddpstep_train_step_indexed)"}}
    if allreduce is not None:
        line["allreduce"] = allreduce
    if large is not None:
        line["large_batch"] = large
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        b = cpu_reference_run(B, steps=3000, warmup=30, budget_s=12.0)
        line["cpu_baseline"] = {k: b[k] for k in ("value", "unit", "cores", "kind", "sample")}
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        print(json.dumps(line), flush=True)


def roofline_of(prof, B, es, hbm_peak, tf_peak, peak_src, n_prof):
    """Roofline of the dominant kernel from the per-kernel event times of ddpstep_profile_step."""
    # slots of kernels that do not run at this batch come back negative: |value| = the bare interval between two event
    # nodes, i.e. the instrumentation overhead every other interval contains; it is subtracted from the kernel times
    empty = [-v for v in prof.values() if v < 0]
    ev_overhead = sum(empty) / len(empty) if empty else 0.0
    fused_away = [k for k, v in prof.items() if v < 0]
    prof = {k: max(v - ev_overhead, 0.0) for k, v in prof.items() if v >= 0}
    # a kernel that hosts another stage at this batch size also moves that stage's bytes
    hosts = {"w2_tiles_refresh": "conv1_fwd", "bn2_relu_pool": "fc_ce", "bn2_bwd_reduce": "fc_ce",
             "bn2_bwd_apply": "conv2_dgrad", "bn1_bwd_reduce": "conv2_dgrad"}
    work = dict(KERNEL_WORK)
    for k in fused_away:
        h = hosts[k]
        work[h] = tuple(a + b for a, b in zip(work[h], KERNEL_WORK[k]))
    dom = max((k for k in prof if k not in ("step", "conv2_wgrad_reduce", "w2_tiles_refresh")), key=lambda k: prof[k])
    fl, ab, xb = work[dom]
    sec = prof[dom] * 1e-3
    if fl >= 627200 and dom.startswith("conv2"):
        ach = fl * B / sec / 1e12
        roof = {"kernel": dom, "bound": "tensor", "achieved": ach, "peak": tf_peak, "unit": "TFLOP/s", "frac": ach / tf_peak}
    else:
        ach = (ab * es + xb) * B / sec / 1e9
        roof = {"kernel": dom, "bound": "hbm", "achieved": ach, "peak": hbm_peak, "unit": "GB/s", "frac": ach / hbm_peak}
    roof.update({"traffic": None, "peak_source": peak_src, "avg_kernel_us": prof[dom] * 1e3,
                 "how": f"CUDA events between the kernel nodes of the step graph, mean of {n_prof} steps, minus the "
                        f"measured interval of an empty slot ({ev_overhead * 1e3:.2f} us)",
                 "event_overhead_us": round(ev_overhead * 1e3, 2), "stages_fused_into_other_kernels": fused_away,
                 "kernel_us": dict({k: round(v * 1e3, 2) for k, v in prof.items()}, serialised=True),
                 "kernel_us_note": "profiling run: side branches serialised onto one stream and programmatic dependent "
                                   "launch off, so these are per-kernel durations for ranking, NOT stage latencies of "
                                   "the hot path (their sum exceeds the step time)"})
    return roof


def large_batch_run(torch, dist, Engine, st, rank, world, local, barrier, max_over_ranks, hbm_peak, tf_peak, peak_src, ALIGN):
    B2, K2, W2, NB2 = 4096, 20, 5, 12  # 12 x 4096 images = 154 MB fp32 per rank: larger than L2
    g = torch.Generator().manual_seed(3407 + rank)
    x2 = (torch.randint(0, 256, (B2 * NB2, 1, 28, 28), dtype=torch.uint8, generator=g).float() / 255).cuda()
    y2 = torch.randint(0, 10, (B2 * NB2,), generator=g).cuda()
    e2 = Engine(dtype="bf16", batch_cap=B2, rank=rank, world=world, device=local)
    e2.connect()
    bn = torch.cat([st.running_mean[0], st.running_var[0], st.running_mean[1], st.running_var[1]])
    e2.load_state(st.flat_params().numpy(), bn.numpy())
    bt = lambda i: (x2[(i % NB2) * B2:(i % NB2 + 1) * B2], y2[(i % NB2) * B2:(i % NB2 + 1) * B2])
    for i in range(W2):
        e2.train_step(*bt(i))
    barrier()
    a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for i in range(ALIGN):
        e2.train_step(*bt(W2 + i))
    a0.record()
    for i in range(K2):
        e2.train_step(*bt(W2 + ALIGN + i))
    a1.record()
    barrier()
    ms = max_over_ranks(a0.elapsed_time(a1)) / K2
    e2.check_device()
    prof = {}
    for i in range(5):
        for k, v in e2.profile_step(*bt(i), in_graph=True).items():
            prof[k] = prof.get(k, 0.0) + v / 5
    barrier()
    roof = roofline_of(prof, B2, 2, hbm_peak, tf_peak, peak_src, 5)
    step_bytes = STEP_BYTES_PER_IMG_AMP * B2 + STEP_PARAM_BYTES
    finite = bool(torch.isfinite(e2.params).all().item())
    scale_ok = (e2.scale.item(), e2.growth_tracker.item()) == (65536.0, W2 + ALIGN + K2 + 5)
    e2.close()
    return {"workload": "BASELINE.json configs[3]: ConvNet AMP bf16, batch 4096/GPU", "value": world * B2 / (ms * 1e-3),
            "unit": UNIT, "ms_per_step": ms, "steps": K2, "warmup": W2 + ALIGN, "dtype": "bf16 storage, f32 accumulate",
            "step_hbm_frac": step_bytes / (ms * 1e-3) / 1e9 / hbm_peak,
            "step_hbm_floor_ms": step_bytes / (hbm_peak * 1e9) * 1e3,
            "step_tflops": STEP_FLOPS_PER_IMG * B2 / (ms * 1e-3) / 1e12,
            "roofline": roof, "params_finite": finite, "scaler_state_ok": scale_ok}


if __name__ == "__main__":
    main()
