"""The reference's OWN scripts, unmodified except for the two import lines of INTEGRATION.md section 1, run on the fused
step -- and the same scripts run untouched on stock torch, from the same seed and the same synthetic idx files.

north_star: "ddp_main.py and ddp_main_torchrun.py run with import-level changes only".  The scripts are staged by
tools/stage_reference.py (build container; /root/reference does not exist on the GPU box) into the git-ignored
baseline/_ref/; without that directory these tests skip.

  * ddp_main.py            -- mp.spawn + hand-set MASTER_ADDR / MASTER_PORT / WORLD_SIZE / RANK (ddp_main.py:61-73,176)
  * ddp_main_torchrun.py   -- torchrun, LOCAL_RANK (ddp_main_torchrun.py:101-104)

Compared after one epoch (12,800 synthetic images, batch 32 per rank, ragged last batch included):
the saved `ddp_checkpoint.pt` -- same keys (`module.*`, 5 scaler keys), scaler dict EQUAL (scale, growth tracker:
bit-exact decisions over the whole epoch), parameter update within tolerance -- and the printed test() accuracy."""
import json
import os
import re
import shutil
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
STAGED = os.path.join(ROOT, "baseline", "_ref")
OUTDIR = os.path.join(ROOT, "gpurun_out", "refrun")


def _need_staged():
    if not os.path.exists(os.path.join(STAGED, "shim_ddp_torchrun.py")):
        pytest.skip("reference scripts not staged (python tools/stage_reference.py in the build container)")


def _run(kind, style, world, port):
    work = os.path.join(OUTDIR, f"{kind}_{style}_w{world}")
    shutil.rmtree(work, ignore_errors=True)
    os.makedirs(work)
    os.symlink(os.path.join(STAGED, "data"), os.path.join(work, "data"))
    script = os.path.join(STAGED, f"{kind}_ddp_{style}.py")
    gpus = ",".join(str(i) for i in range(world))
    env = dict(os.environ, PYTHONPATH=os.pathsep.join([os.path.join(STAGED, "seed"), ROOT]))
    env.pop("CUDA_VISIBLE_DEVICES", None)
    if style == "spawn":
        cmd = [sys.executable, script, "--gpu", gpus, "-e", "1", "-b", "32"]
    else:
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
               "--master-addr", "127.0.0.1", "--master-port", str(port), script, "--gpu", gpus, "-e", "1", "-b", "32"]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=work, env=env)
    open(os.path.join(work, "stdout.log"), "w").write(res.stdout + "\n--- stderr ---\n" + res.stderr[-4000:])
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-3000:]
    acc = float(re.search(r"Accuracy is ([0-9.]+)%", res.stdout).group(1))
    ckpt = torch.load(os.path.join(work, "ddp_checkpoint.pt"), map_location="cpu", weights_only=False)
    return acc, ckpt, res.stdout


def _compare(style, world, port):
    _need_staged()
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    acc_s, ck_s, out_s = _run("shim", style, world, port)
    acc_r, ck_r, out_r = _run("ref", style, world, port + 1)
    assert list(ck_s["model"].keys()) == list(ck_r["model"].keys())
    assert all(k.startswith("module.") for k in ck_s["model"])
    assert ck_s["scaler"] == ck_r["scaler"], (ck_s["scaler"], ck_r["scaler"])
    # same init on both sides (seeded at interpreter start-up): rebuild it to measure the UPDATE, not the weights
    from tests.refmodel import RefConvNet
    torch.manual_seed(3407)
    init = {"module." + k: v for k, v in RefConvNet(None, "cpu").state_dict().items()}
    worst = 0.0
    for k, v in ck_r["model"].items():
        if "num_batches_tracked" in k:
            assert int(ck_s["model"][k]) == int(v)
            continue
        a, b = ck_s["model"][k].float(), v.float()
        if "running" in k:
            assert torch.allclose(a, b, rtol=5e-3, atol=1e-5), k
            continue
        # a conv bias feeds straight into a BatchNorm: its gradient is mathematically zero and its "update" is rounding
        # noise on both sides -- measure it on the scale of the sibling weight's update (as tests/helpers.py does)
        kk = k.replace(".0.bias", ".0.weight") if k.endswith(".0.bias") else k
        upd_ref = (ck_r["model"][kk].float() - init[kk]).norm().item()
        err = (a - b).norm().item() / max(upd_ref, 1e-12)
        worst = max(worst, err)
        assert upd_ref > 0 and err < 0.1, (k, err)
    assert abs(acc_s - acc_r) <= 1.0, (acc_s, acc_r)
    rec = {"style": style, "world": world, "accuracy_shim": acc_s, "accuracy_torch": acc_r, "scaler": ck_s["scaler"],
           "worst_update_rel_err": worst}
    with open(os.path.join(OUTDIR, "summary.jsonl"), "a") as f:
        f.write(json.dumps(rec) + "\n")
    return rec


def test_ddp_main_mp_spawn_script_runs_on_the_fused_step_w1():
    """ddp_main.py as written (mp.spawn, its own env bootstrap), `--gpu 0`: world size 1, SyncBN falls back to BN."""
    rec = _compare("spawn", 1, 0)
    assert rec["scaler"]["scale"] == 32768.0 and rec["scaler"]["_growth_tracker"] == 399  # overflow at step 0, then 399 good steps


def test_ddp_main_torchrun_script_runs_on_the_fused_step_all_gpus():
    """ddp_main_torchrun.py under torchrun on every GPU of the box (2, 4 or 8): SyncBN + DDP all-reduce per step."""
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    world = 8 if n >= 8 else (4 if n >= 4 else 2)
    rec = _compare("torchrun", world, 29811)
    steps = -(-(12800 // world) // 32)
    assert rec["scaler"]["scale"] == 32768.0 and rec["scaler"]["_growth_tracker"] == steps - 1


def test_ddp_main_mp_spawn_script_two_gpus():
    rec = _compare("spawn", 2, 0)
    assert rec["scaler"]["_growth_tracker"] == 199
