"""bench.py contract pieces that run without a GPU: the reference arm (the UNMODIFIED reference over the
gloo mpi4py shim) and the `unavailable` answer for configs the reference cannot run."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "baseline", "_ref", "dataParallelTraining_NN_MPI.py")


def _run(args, nproc=1, timeout=300):
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", OMP_NUM_THREADS="1")
    if nproc == 1:
        cmd = [sys.executable, os.path.join(ROOT, "bench.py"), *args]
    else:
        import socket
        s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
               "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "bench.py"), *args]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1, r.stdout
    return json.loads(lines[0])


def test_reference_arm_reports_unavailable_for_the_wide_config():
    d = _run(["--impl", "reference"])
    assert d["impl"] == "reference" and "unavailable" in d and "2-3-1" in d["unavailable"]


@pytest.mark.skipif(not os.path.exists(REF), reason="baseline/_ref not installed (baseline/install_reference.sh)")
def test_reference_arm_runs_the_unmodified_reference():
    upstream = "/root/reference/dataParallelTraining_NN_MPI.py"
    if os.path.exists(upstream):
        with open(REF, "rb") as a, open(upstream, "rb") as b:
            assert a.read() == b.read()                      # byte-identical to the reference
    d = _run(["--impl", "reference", "--config", "ref", "--steps", "5", "--warmup", "3"])
    assert d["impl"] == "reference" and d["n_gpus"] == 1 and d["steps"] == 5
    assert d["value"] > 0 and d["ms_per_step"] > 0 and d["config"]["global_batch"] == 16


@pytest.mark.skipif(not os.path.exists(REF), reason="baseline/_ref not installed")
@pytest.mark.timeout(600)
def test_reference_arm_two_ranks_and_uneven_world():
    d = _run(["--impl", "reference", "--config", "ref", "--gpus", "2", "--steps", "4", "--warmup", "3"], nproc=2)
    assert d["n_gpus"] == 2 and d["value"] > 0
    u = _run(["--impl", "reference", "--config", "ref", "--gpus", "3", "--steps", "4", "--warmup", "3"], nproc=3)
    assert "unavailable" in u and "MPI.INT" in u["unavailable"]
