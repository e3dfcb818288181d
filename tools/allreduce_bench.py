#!/usr/bin/env python
"""The DDP gradient all-reduce in isolation (north_star: "allreduce latency and bus GB/s against NVLink peak").

    torchrun --nproc-per-node W tools/allreduce_bench.py [--iters 1000] [--out gpurun_out/allreduce_wW.json]

The reference's collective is ONE ncclAllReduce of the 29,034-element fp32 bucket per step (ddp_main.py:121-123,
reducer.hpp:524), followed by ~8 more kernels for unscale / inf check / SGD / scale update (grad_scaler.py:294-298,
sgd.py:473).  Here all of that is one kernel (k_step); `ddpstep_allreduce_only` launches it alone.  Compared, on the
same box and the same message size, device-timed with CUDA events around back-to-back launches (max over ranks):

  ours            ddpstep_allreduce_only: exchange + unscale + non-finite check + SGD + scale update
  nccl            dist.all_reduce(bucket)                      -- what the reference calls (NCCL 2.28.9)
  nccl_graph      the same, 50 calls captured in one CUDA graph (no host launch overhead)
  symm_one_shot   torch.ops.symm_mem.one_shot_all_reduce       -- torch's own P2P one-shot kernel
  symm_multimem   torch.ops.symm_mem.multimem_one_shot_all_reduce (NVLS in-switch reduction), if the box allows it
  nccl_small      all_gather of 33 floats / all_reduce of 64 floats: the reference's SyncBN exchanges (C3-C6)

busbw = algbw * 2 (W - 1) / W (NCCL's convention), against 900 GB/s per direction of NVLink 5."""
import argparse
import json
import os
import statistics
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from ddp_practice_b200 import Engine  # noqa: E402
from oracle import convnet_oracle as O  # noqa: E402  (synthetic inputs / random-init weights only)
from oracle.reference_step import build_model  # noqa: E402

S_BYTES = 116136


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=1000)
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v):
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item()

    def timed(fn, iters=args.iters, per_call=1):
        """mean us per call of `iters` back-to-back calls, and the p50 of per-call event intervals"""
        for _ in range(20):
            fn()
        barrier()
        for _ in range(3):
            fn()
        evs = [torch.cuda.Event(enable_timing=True) for _ in range(iters + 1)]
        evs[0].record()
        for i in range(iters):
            fn()
            evs[i + 1].record()
        barrier()
        mean = evs[0].elapsed_time(evs[-1]) / iters * 1e3 / per_call
        p50 = statistics.median(evs[i].elapsed_time(evs[i + 1]) for i in range(iters)) * 1e3 / per_call
        return {"mean_us": max_over_ranks(mean), "p50_us": max_over_ranks(p50)}

    def bw(us):
        alg = S_BYTES / (us * 1e-6) / 1e9
        return {"algbw_gbs": alg, "busbw_gbs": alg * 2 * (world - 1) / world}

    res = {"world": world, "bytes": S_BYTES, "iters": args.iters, "nvlink_peak_gbs_per_dir": 900.0,
           "nccl_version": ".".join(map(str, torch.cuda.nccl.version())), "torch": torch.__version__}

    # ---- ours
    st = O.OracleState.from_state_dict(build_model().state_dict())
    eng = Engine(dtype="fp16", batch_cap=32, rank=rank, world=world, device=local)
    eng.connect()
    eng.load_state(st.flat_params().numpy(), torch.cat([st.running_mean[0], st.running_var[0], st.running_mean[1], st.running_var[1]]).numpy())
    x, y = O.synthetic_mnist(64 * world)
    x, y = x[rank::world].cuda(), y[rank::world].cuda()
    for s in range(2):
        eng.train_step(x[s * 32:(s + 1) * 32].contiguous(), y[s * 32:(s + 1) * 32].contiguous())
    torch.cuda.synchronize()
    r = timed(eng.allreduce_only)
    res["ours_fused_tail_eager"] = dict(r, **bw(r["p50_us"]), what="ddpstep_allreduce_only, eager launches from Python (host-bound): "
                                        "all-reduce + unscale + non-finite check + SGD + scale update")
    eng.check_device()

    def graphed(fn, what, key):
        """50 calls per CUDA-graph replay: no host launch overhead, the device-side cost of one call"""
        try:
            g = torch.cuda.CUDAGraph()
            sg = torch.cuda.Stream()
            with torch.cuda.stream(sg):
                for _ in range(3):
                    fn()
                torch.cuda.synchronize()
                with torch.cuda.graph(g, stream=sg):
                    for _ in range(50):
                        fn()
            r = timed(g.replay, iters=max(args.iters // 50, 10), per_call=50)
            res[key] = dict(r, **bw(r["p50_us"]), what=what + "; 50 calls per CUDA-graph replay")
        except Exception as e:  # noqa: BLE001
            res[key] = {"unavailable": repr(e)[:200]}

    graphed(eng.allreduce_only, "the fused tail (all-reduce + unscale + inf check + SGD + scale update)", "ours_fused_tail")
    eng.check_device()
    plain = torch.randn(29034, device="cuda")
    graphed(lambda: eng.allreduce_bucket(plain), "ddpstep_allreduce_bucket: the bare exchange of the same 116,136 bytes "
            "(the comm-hook entry point)", "ours_bare_exchange")
    eng.check_device()

    # ---- NCCL, as the reference calls it
    bucket = torch.randn(29034, device="cuda")
    r = timed(lambda: dist.all_reduce(bucket))
    res["nccl"] = dict(r, **bw(r["p50_us"]), what="dist.all_reduce, eager launches")
    graphed(lambda: dist.all_reduce(bucket), "dist.all_reduce (NCCL)", "nccl_graph")

    # ---- torch symmetric memory one-shot kernels
    try:
        import torch.distributed._symmetric_memory as symm_mem
        t = symm_mem.empty(29034, dtype=torch.float32, device=f"cuda:{local}")
        symm_mem.rendezvous(t, dist.group.WORLD.group_name)
        t.normal_()
        name = dist.group.WORLD.group_name
        r = timed(lambda: torch.ops.symm_mem.one_shot_all_reduce(t, "sum", name))
        res["symm_one_shot"] = dict(r, **bw(r["p50_us"]), what="torch.ops.symm_mem.one_shot_all_reduce")
        graphed(lambda: torch.ops.symm_mem.one_shot_all_reduce(t, "sum", name), "torch.ops.symm_mem.one_shot_all_reduce", "symm_one_shot_graph")
        try:
            r = timed(lambda: torch.ops.symm_mem.multimem_one_shot_all_reduce(t, "sum", name))
            res["symm_multimem"] = dict(r, **bw(r["p50_us"]), what="torch.ops.symm_mem.multimem_one_shot_all_reduce (NVLS)")
            graphed(lambda: torch.ops.symm_mem.multimem_one_shot_all_reduce(t, "sum", name), "multimem_one_shot_all_reduce (NVLS)", "symm_multimem_graph")
        except Exception as e:  # noqa: BLE001
            res["symm_multimem"] = {"unavailable": repr(e)[:200]}
    except Exception as e:  # noqa: BLE001
        res["symm_one_shot"] = {"unavailable": repr(e)[:200]}

    # ---- the reference's SyncBN-sized collectives (C3-C6 of SURVEY.md 2.3)
    small_in = torch.randn(33, device="cuda")
    small_out = torch.empty(33 * world, device="cuda")
    r = timed(lambda: dist.all_gather_into_tensor(small_out, small_in))
    res["nccl_allgather_132B"] = r
    small = torch.randn(64, device="cuda")
    r = timed(lambda: dist.all_reduce(small))
    res["nccl_allreduce_256B"] = r

    # (no eng.close(): the process leaves through os._exit below)
    if rank == 0:
        line = json.dumps(res)
        print(line, flush=True)
        if args.out:
            os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
            with open(args.out, "w") as f:
                f.write(line + "\n")
    # CUDA graphs holding NCCL / symmetric-memory work make an orderly teardown hang on this stack: results are out, leave
    sys.stdout.flush()
    torch.cuda.synchronize()
    os._exit(0)


if __name__ == "__main__":
    main()
