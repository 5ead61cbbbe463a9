"""BASELINE config 5's exchange on real GPUs: spawns tests/multi_gpu_check.py under torch.distributed.run with one process
per GPU (2 GPUs, and all of them when the box has more) and demands that every rank reports parity with the oracle's
lists of the unsharded scene.  Skipped on a one-GPU box, where tests/test_parity_gpu.py runs the same protocol with the
shards sharing the GPU (test_group_exchange_gathers_the_unsharded_lists) and tests/test_shard_gloo.py its host logic."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _gpus():
    from pmtech_b200 import capi
    return capi.device_count()


@pytest.mark.gpu
@pytest.mark.parametrize("procs", [2, 8])
def test_exchange_across_gpus_matches_the_oracle(gpu, oracle, procs):
    have = _gpus()
    if have < 2:
        pytest.skip("one GPU: the multi-process exchange needs two")
    if procs > have:
        if procs == 8 and have > 2:
            procs = have
        else:
            pytest.skip(f"{have} GPUs on this box")
    env = dict(os.environ, OMP_NUM_THREADS="4")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={procs}", "--master-addr", "127.0.0.1",
           "--master-port", str(29533 + procs), os.path.join(ROOT, "tests", "multi_gpu_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=420, cwd=ROOT, env=env)
    out_dir = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out_dir, exist_ok=True)
    with open(os.path.join(out_dir, f"multi_gpu_check_n{procs}.log"), "w") as fh:
        fh.write(r.stdout + "\n--- stderr ---\n" + r.stderr[-4000:])
    assert r.returncode == 0 and "EXCHANGE CHECK PASSED" in r.stdout, (r.stdout[-3000:], r.stderr[-3000:])
