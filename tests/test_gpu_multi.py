"""Real multi-GPU runs (one process per GPU over CUDA-IPC symmetric memory): P6-P9, P25."""
import json
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpu():
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


@pytest.mark.skipif(_ngpu() < 2, reason="needs >= 2 GPUs")
def test_multi_gpu_trainers(native_lib, tmp_path):
    n = min(_ngpu(), 8)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
           "--master-port", "29613", os.path.join(ROOT, "tests", "multi_gpu_worker.py"), str(tmp_path)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    if os.path.isdir(os.path.join(ROOT, "gpurun_out")):          # keep the ranks' output for post-mortems
        open(os.path.join(ROOT, "gpurun_out", "multi_worker.log"), "w").write(r.stdout + r.stderr)
    assert r.returncode == 0, (r.stdout + r.stderr)[-6000:]
    res = json.load(open(tmp_path / "result.json"))
    assert res["world"] == n
    assert res["sync_digests_equal"] and res["sync_global_step"] == 60
    assert res["sync_loss"][1] < res["sync_loss"][0] and res["sync_acc"] > 0.2
    # the push ran inside the step kernel and reproduced the separate-kernel push bit for bit; the tf32
    # tensor-core chain lands within tf32 rounding of the exact-fp32 fused step
    assert res["sync_fused_push_used"] and res["sync_fused_push_bitwise"], res
    assert res["sync_chain_close"] < 5e-3 and res["sync_steps_all"] == [60, 60], res
    assert res["async_global_step"] == 40 * n and res["async_loss"][1] < res["async_loss"][0]
    assert res["async_in_kernel"]              # the Hogwild step ran as one launch against the owners' masters
    assert res["wide_digests_equal"] and res["wide_loss"][1] < res["wide_loss"][0]
    assert res["wide_fused_pull"] == (512 % (64 * n) == 0)
    assert res["sparse_async_global_step"] == 10 * n and res["sparse_sync_global_step"] == 10
    assert res["barrier_epoch"] == 5
    # a checkpoint taken on n GPUs restores on one (P25)
    from b200ps.io.saver import Saver, latest_checkpoint
    from b200ps.parallel.context import init_from_env
    from b200ps.train import MLPTrainer
    from b200ps import ops
    ck = latest_checkpoint(str(tmp_path))
    assert ck.endswith("model.ckpt-60")
    ctx = init_from_env(virtual_world=1)
    tr = MLPTrainer(ctx, batch_size=100, lr=0.01, mode="sync", use_graph=False)
    tr.store.load_state_dict(Saver.restore(ck))
    assert tr.global_step == 60
    x = torch.empty(2000, 784, device="cuda"); y = torch.empty(2000, device="cuda", dtype=torch.int32)
    ops.synth_mnist(x, y, seed=1, offset=0)
    acc, _ = tr.evaluate(x, y)
    assert abs(acc - res["sync_acc"]) < 1e-6


@pytest.mark.skipif(_ngpu() < 2, reason="needs >= 2 GPUs")
def test_nvls_multimem_push(native_lib, tmp_path):
    """K1-NVLS: multimem.ld_reduce / multimem.st push on torch symmetric memory == mean-gradient SGD on every replica."""
    n = min(_ngpu(), 8)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
           "--master-port", "29617", os.path.join(ROOT, "tests", "nvls_worker.py"), str(tmp_path)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    if os.path.isdir(os.path.join(ROOT, "gpurun_out")):
        open(os.path.join(ROOT, "gpurun_out", "nvls_worker.log"), "w").write(r.stdout + r.stderr)
    assert r.returncode == 0, (r.stdout + r.stderr)[-4000:]
    res = json.load(open(tmp_path / "nvls_result.json"))
    if "error" in res and "multicast is not available" in res["error"]:
        pytest.skip("no NVLS multicast on this system")
    assert "error" not in res, res
    assert res["ok"] and res["replicas_identical"] and res["global_step"] == 3, res


def _torchrun(n, script, args, port, timeout=600):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "tests", script)] + [str(a) for a in args]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout)
    if os.path.isdir(os.path.join(ROOT, "gpurun_out")):
        open(os.path.join(ROOT, "gpurun_out", script.replace(".py", f"_n{n}.log")), "w").write(r.stdout + r.stderr)
    return r


@pytest.mark.skipif(_ngpu() < 2, reason="needs >= 2 GPUs")
def test_multi_gpu_results_match_independent_oracles(native_lib, tmp_path):
    """Sync MLP == float64 big-batch SGD; replicas_to_aggregate < world == the documented rotating set; async pushes
    conserve every update; sparse sync weight VALUES == the float64 dense restatement (oracles: plain PyTorch)."""
    n = min(_ngpu(), 8)
    r = _torchrun(n, "multi_gpu_oracle_worker.py", [tmp_path], 29621)
    assert r.returncode == 0, (r.stdout + r.stderr)[-6000:]
    res = json.load(open(tmp_path / "oracle_result.json"))
    assert res["world"] == n
    assert res["sync_mlp_max_abs_err"] < 2e-4 and res["sync_mlp_global_step"] == 12, res
    assert res["sync_backup_max_abs_err"] < 2e-4 and res["sync_backup_global_step"] == 12, res
    assert res["async_conserved"] and res["async_global_step"] == 25 * n, res
    assert res["sparse_sync_max_abs_err"] < 1e-4 and res["sparse_sync_bias_err"] < 1e-5, res
    assert res["sparse_sync_global_step"] == 5, res
    # the unique-id step on ids that repeat (the reference's gather-once / update-once plan, lr2.py:390-396)
    assert res["sparse_unique_sync_max_abs_err"] < 1e-4 and res["sparse_unique_sync_bias_err"] < 1e-5, res
    assert res["sparse_unique_sync_global_step"] == 5 and res["sparse_unique_distinct_ratio"] < 0.8, res
    # large groups: through the switch when the box has NVLS (else the same numbers from the P2P kernels)
    assert res["nvls_f32_master_err"] < 1e-5 and res["nvls_f32_work_err"] < 1e-5 and res["nvls_f32_pull_err"] < 1e-5, res
    assert res["nvls_bf16_master_err"] < 2e-2 and res["nvls_bf16_work_err"] < 2e-2 and res["nvls_bf16_pull_err"] < 2e-2, res
    assert res["nvls_f32_replicas_identical"] and res["nvls_bf16_replicas_identical"], res
    assert res["nvls_f32_global_step"] == 3 and res["nvls_bf16_global_step"] == 3, res
    # wide & deep, synchronous, owner-side rows: no embedding row crossed a link, the tables match the oracle
    assert res["wd_owner_step"] and res["wd_sync_global_step"] == 3, res
    assert max(res["wd_sync_E_err"], res["wd_sync_W_err"], res["wd_sync_b_err"]) < 2e-5 and res["wd_sync_w2_err"] < 1e-4, res


@pytest.mark.skipif(_ngpu() < 2, reason="needs >= 2 GPUs")
def test_nvlink_backend_detects_a_dead_rank(native_lib, tmp_path):
    """A worker (= a parameter-server shard) dies between steps: the survivors raise PeerLost from the liveness
    check within its timeout instead of launching a step that would wait on the dead rank, and exit normally."""
    r = _torchrun(2, "multi_gpu_fault_worker.py", [tmp_path, 7], 29625, timeout=300)
    assert r.returncode == 0, (r.stdout + r.stderr)[-4000:]
    res = json.load(open(tmp_path / "fault_result.json"))
    assert res.get("lost") and res["steps_done"] == 7 and res["global_step"] == 7, res
    assert res["detect_seconds"] < 8.0 and "stopped answering" in res["message"], res


@pytest.mark.skipif(_ngpu() < 2, reason="needs >= 2 GPUs")
def test_sync_push_survives_injected_skew_on_checked_build(native_lib, tmp_path):
    """10 000 synchronous steps with random per-CTA delays on the library built with -DB200PS_CHECKED: no device
    assertion fires, replicas stay identical, and the result equals the undisturbed run bit for bit."""
    import build as _build
    _build.build(quiet=True, checked=True)          # incremental: a no-op when the checked library is current
    n = min(_ngpu(), 8)
    r = _torchrun(n, "multi_gpu_stress_worker.py", [tmp_path, 10000, 20000], 29629, timeout=600)
    assert r.returncode == 0, (r.stdout + r.stderr)[-4000:]
    res = json.load(open(tmp_path / "stress_result.json"))
    assert res["bitwise_equal"] and res["plain"]["replicas_identical"] and res["skewed"]["replicas_identical"], res
    assert res["plain"]["global_step"] == res["skewed"]["global_step"] == 10003, res
