"""BASELINE.json configs[0]: the reference MLP on make_regression data, 2 processes on CPU (gloo)."""
import os
import re
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    import socket
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _torchrun(nproc, script_args, timeout=300):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()),
           os.path.join(ROOT, "train.py"), *script_args]
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", OMP_NUM_THREADS="1")
    return subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, env=env, cwd=ROOT)


def _losses(out, rank):
    return [float(x) for x in re.findall(rf"loss in worker {rank}: ([-0-9.e+nan]+)", out)]


@pytest.mark.timeout(600)
def test_torchrun_two_ranks_reference_config(tmp_path):
    ck = str(tmp_path / "ck.pt")
    r = _torchrun(2, ["--checkpoint", ck, "--check-replicas-every", "1"])
    assert r.returncode == 0, r.stderr[-2000:]
    assert r.stdout.count("[ = = = = = Epoch 0 = = = = = ]") == 2          # every rank prints (Q10)
    assert _losses(r.stdout, 0) == pytest.approx([2230.0781, 2227.2244, 2221.5251], rel=2e-6)
    assert _losses(r.stdout, 1) == pytest.approx([2835.1191, 2831.0088, 2822.7883], rel=2e-6)
    assert os.path.exists(ck)


@pytest.mark.timeout(600)
def test_torchrun_three_ranks_uneven_allreduce():
    r = _torchrun(3, ["--sync", "allreduce"])
    assert r.returncode == 0, r.stderr[-2000:]
    assert _losses(r.stdout, 0) == pytest.approx([2448.8511, 2444.9055, 2437.0879], rel=2e-6)
    assert _losses(r.stdout, 2) == pytest.approx([1471.8096, 1471.6271, 1471.2463], rel=2e-6)


def test_mpiexec_style_env_single_rank():
    env = dict(os.environ, OMPI_COMM_WORLD_RANK="0", OMPI_COMM_WORLD_SIZE="1", CUDA_VISIBLE_DEVICES="")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "dataParallelTraining_NN_MPI.py"), "--nepochs", "1"],
                       capture_output=True, text=True, env=env, cwd=ROOT, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    assert _losses(r.stdout, 0) == pytest.approx([2530.8320], rel=2e-6)


@pytest.mark.timeout(600)
def test_mpiexec_style_launch_two_ranks():
    """What `mpiexec -n 2 python dataParallelTraining_NN_MPI.py` does (README:12): every process gets
    only Open MPI's rank variables; the trainer must find its rank and peers from them."""
    port = _free_port()
    procs = []
    for r in range(2):
        env = dict(os.environ, OMPI_COMM_WORLD_RANK=str(r), OMPI_COMM_WORLD_SIZE="2",
                   OMPI_COMM_WORLD_LOCAL_RANK=str(r), MASTER_PORT=str(port), CUDA_VISIBLE_DEVICES="",
                   OMP_NUM_THREADS="1")
        for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
            env.pop(k, None)
        procs.append(subprocess.Popen([sys.executable, os.path.join(ROOT, "dataParallelTraining_NN_MPI.py")],
                                      stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, env=env, cwd=ROOT))
    outs = [p.communicate(timeout=500) for p in procs]
    assert all(p.returncode == 0 for p in procs), [o[1][-1500:] for o in outs]
    assert _losses(outs[0][0], 0) == pytest.approx([2230.0781, 2227.2244, 2221.5251], rel=2e-6)
    assert _losses(outs[1][0], 1) == pytest.approx([2835.1191, 2831.0088, 2822.7883], rel=2e-6)
