"""Launcher contract: rank / world discovery and device binding.

The reference is started with ``mpiexec -n N python script.py`` (README:12) and reads its rank
from ``MPI.COMM_WORLD`` (ref.py:61-63).  Here a process learns who it is from the environment, in
this order: torchrun (``RANK/WORLD_SIZE/LOCAL_RANK``), Open MPI (``OMPI_COMM_WORLD_*``), MPICH /
Intel MPI / Slurm PMI (``PMI_RANK/PMI_SIZE``, ``PMIX_RANK``, ``SLURM_PROCID``).  No variables means a
single process.  One process drives one GPU (``cudaSetDevice(LOCAL_RANK)``).
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import List, Mapping, Optional


@dataclass(frozen=True)
class RankInfo:
    rank: int
    world_size: int
    local_rank: int
    launcher: str
    master_addr: str = "127.0.0.1"
    master_port: int = 29500


def _first_int(env: Mapping[str, str], *names: str) -> Optional[int]:
    for n in names:
        v = env.get(n)
        if v not in (None, ""):
            return int(v)
    return None


def discover(env: Optional[Mapping[str, str]] = None) -> RankInfo:
    env = os.environ if env is None else env
    addr = env.get("MASTER_ADDR", "127.0.0.1")
    port = int(env.get("MASTER_PORT", "29500"))
    if "RANK" in env and "WORLD_SIZE" in env:
        rank, world = int(env["RANK"]), int(env["WORLD_SIZE"])
        local = _first_int(env, "LOCAL_RANK")
        return RankInfo(rank, world, rank if local is None else local, "torchrun", addr, port)
    if "OMPI_COMM_WORLD_RANK" in env:
        rank = int(env["OMPI_COMM_WORLD_RANK"])
        world = int(env.get("OMPI_COMM_WORLD_SIZE", "1"))
        local = _first_int(env, "OMPI_COMM_WORLD_LOCAL_RANK")
        return RankInfo(rank, world, rank if local is None else local, "openmpi", addr, port)
    rank = _first_int(env, "PMI_RANK", "PMIX_RANK", "SLURM_PROCID")
    if rank is not None:
        world = _first_int(env, "PMI_SIZE", "SLURM_NTASKS") or 1
        local = _first_int(env, "MPI_LOCALRANKID", "SLURM_LOCALID")
        return RankInfo(rank, world, rank if local is None else local, "pmi", addr, port)
    return RankInfo(0, 1, 0, "single", addr, port)


def pcie_closeness() -> Optional[List[List[int]]]:
    """``c[i][j]`` = how much of the PCIe path to the host the visible GPUs i and j share (larger =
    closer = more contention for host copies).  NVLink says nothing about this: on an HGX board every
    pair is NV18, yet pairs of GPUs hang off one PCIe switch and share its x16 uplink - two ranks that
    feed batches from host memory through the same uplink each get half of it.  Source: NVML's common
    ancestor level (PIX/PXB = same switch), else the sysfs device paths; ``None`` if neither is
    available."""
    try:
        import pynvml
        import torch
        n = torch.cuda.device_count()
        pynvml.nvmlInit()
        hs = [pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + str(torch.cuda.get_device_properties(i).uuid)).encode())
              for i in range(n)]
    except Exception:
        return None
    score = [[0] * n for _ in range(n)]
    ok = False
    try:
        # NVML_TOPOLOGY_INTERNAL 0, SINGLE 10, MULTIPLE 20, HOSTBRIDGE 30, NODE 40, SYSTEM 50
        for i in range(n):
            for j in range(i + 1, n):
                lvl = int(pynvml.nvmlDeviceGetTopologyCommonAncestor(hs[i], hs[j]))
                score[i][j] = score[j][i] = 60 - lvl
        ok = len({score[i][j] for i in range(n) for j in range(n) if i != j}) > 1
    except Exception:
        ok = False
    if not ok:
        try:        # sysfs: number of shared path components below the PCI root
            paths = []
            for h in hs:
                bus = pynvml.nvmlDeviceGetPciInfo(h).busId
                bus = bus.decode() if isinstance(bus, bytes) else bus
                name = bus.lower()[-12:]                      # 00000000:1B:00.0 -> 0000:1b:00.0
                paths.append(os.path.realpath(f"/sys/bus/pci/devices/{name}").split("/"))
            for i in range(n):
                for j in range(i + 1, n):
                    k = 0
                    while k < min(len(paths[i]), len(paths[j])) and paths[i][k] == paths[j][k]:
                        k += 1
                    score[i][j] = score[j][i] = k
            ok = len({score[i][j] for i in range(n) for j in range(n) if i != j}) > 1
        except Exception:
            ok = False
    return score if ok else None


_PROBE_SRC = r"""
import json, sys, torch
count = int(sys.argv[1])
n = torch.cuda.device_count()
nbytes = 48 << 20
host, dev, st = [], [], []
for i in range(n):
    torch.cuda.set_device(i)
    host.append(torch.empty(nbytes, dtype=torch.uint8).pin_memory())
    dev.append(torch.empty(nbytes, dtype=torch.uint8, device=f"cuda:{i}"))
    st.append(torch.cuda.Stream(device=i))
def run(ids, reps=3):
    # every GPU in `ids` copies 2 x 48 MB from pinned host memory at the same time; worst GPU's best time
    best = {i: 1e9 for i in ids}
    for _ in range(reps):
        ev = {}
        for i in ids:
            torch.cuda.set_device(i)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            with torch.cuda.stream(st[i]):
                a.record(st[i]); dev[i].copy_(host[i], non_blocking=True); dev[i].copy_(host[i], non_blocking=True); b.record(st[i])
            ev[i] = (a, b)
        for i in ids:
            torch.cuda.set_device(i); st[i].synchronize()
            best[i] = min(best[i], ev[i][0].elapsed_time(ev[i][1]))
    return max(best.values())
for i in range(n):
    run([i], 1)
alone = [run([i]) for i in range(n)]
chosen, log = [0], []
while len(chosen) < min(count, n):
    cands = [(run(chosen + [g]), g) for g in range(n) if g not in chosen]
    t, g = min(cands)
    log.append({"add": g, "ms": t, "all": [[round(a, 3), b] for a, b in cands]})
    chosen.append(g)
json.dump({"devices": chosen, "alone_ms": alone, "group_ms": run(chosen), "first_n_ms": run(list(range(min(count, n)))),
           "log": log}, sys.stdout)
"""


def measured_spread(count: int, timeout_s: float = 120.0) -> Optional[List[int]]:
    """Where the host hides its PCIe / NUMA layout (virtualised boxes give every GPU its own root complex,
    NVML reports every pair as "same node", there is one NUMA node), pick the ``count`` GPUs by
    measurement: start from GPU 0 and keep adding the GPU with which the group's simultaneous pinned
    H2D copies finish soonest.  On this pool's 8-GPU boxes pairs never slow each other down, but four
    GPUs of the same half of the box share ~115 GB/s of host bandwidth, so GPUs 0-3 feed at half rate
    while two GPUs from each half feed at full rate.  Runs once per job, in a short-lived child process of
    local rank 0 (no rank keeps CUDA contexts on GPUs it does not use); the other local ranks read the
    result from a file keyed by the launcher's pid.  Returns None when the probe fails or finds nothing
    better than GPUs 0..count-1."""
    import json
    import subprocess
    import sys
    import tempfile
    import time

    local_rank = _first_int(os.environ, "LOCAL_RANK", "OMPI_COMM_WORLD_LOCAL_RANK", "MPI_LOCALRANKID", "SLURM_LOCALID") or 0
    # the launcher process (torchrun agent / mpiexec / shell) is the same for every local rank; its pid plus
    # its start time is unique per launch, so a stale file of an earlier job can never be picked up
    ppid = os.getppid()
    try:
        with open(f"/proc/{ppid}/stat") as f:
            born = f.read().rsplit(")", 1)[1].split()[19]
    except Exception:
        born = "0"
    key = f"b200dp_pcie_{ppid}_{born}_{os.environ.get('MASTER_PORT', '0')}_{count}.json"
    path = os.path.join(tempfile.gettempdir(), key)
    if local_rank == 0:
        result = None
        try:
            r = subprocess.run([sys.executable, "-c", _PROBE_SRC, str(count)], capture_output=True, text=True,
                               timeout=timeout_s)
            if r.returncode == 0 and r.stdout.strip().startswith("{"):
                result = json.loads(r.stdout.strip())
        except Exception:
            result = None
        tmp = path + f".{os.getpid()}.tmp"
        with open(tmp, "w") as f:
            json.dump(result or {"devices": None}, f)
        os.replace(tmp, path)
    else:
        t0 = time.time()
        while not os.path.exists(path) and time.time() - t0 < timeout_s + 30:
            time.sleep(0.05)
    try:
        with open(path) as f:
            res = json.load(f)
    except Exception:
        return None
    devs = res.get("devices")
    if not devs or len(devs) != count or len(set(devs)) != count:
        return None
    if res.get("group_ms", 0) > 0.93 * res.get("first_n_ms", 0):      # nothing to gain: keep 0..count-1
        return None
    return [int(d) for d in devs]


def spread_devices(count: int, closeness: Optional[List[List[int]]], n_visible: int) -> List[int]:
    """``count`` of the ``n_visible`` GPUs, chosen greedily so that the closest pair among them is as
    far apart on PCIe as possible (one GPU per switch before any switch gets a second one).
    Deterministic: every rank computes the same list.  Without topology data: 0, 1, 2, ..."""
    count = min(count, n_visible)
    if closeness is None or count >= n_visible:
        return list(range(count))
    chosen = [0]
    while len(chosen) < count:
        best, best_key = None, None
        for d in range(n_visible):
            if d in chosen:
                continue
            key = (max(closeness[d][c] for c in chosen), sum(closeness[d][c] for c in chosen), d)
            if best_key is None or key < best_key:
                best, best_key = d, key
        chosen.append(best)
    return chosen


def device_index_for(local_rank: int, local_world: Optional[int] = None) -> int:
    """GPU index for this rank.  When the job uses fewer GPUs than are visible (a 2- or 4-rank run on
    an 8-GPU box) the ranks are spread over distinct PCIe uplinks instead of taking GPUs 0..N-1.  The
    uplink sharing comes from NVML / sysfs, or - where the host hides its PCIe tree - from a one-off
    measurement (``measured_spread``; ``B200DP_PCIE_PROBE=0`` switches it off)."""
    import torch

    n = torch.cuda.device_count()
    if local_world is None:
        local_world = _first_int(os.environ, "LOCAL_WORLD_SIZE", "OMPI_COMM_WORLD_LOCAL_SIZE", "WORLD_SIZE") or 1
    if os.environ.get("B200DP_PLACEMENT", "spread") != "spread" or local_world >= n or local_world <= 1:
        return local_rank % max(n, 1)
    closeness = pcie_closeness()
    devs = None
    if closeness is None and os.environ.get("B200DP_PCIE_PROBE", "1") != "0":
        devs = measured_spread(local_world)
    if devs is None:
        devs = spread_devices(local_world, closeness, n)
    return devs[local_rank % len(devs)]


def bind_device(info: RankInfo):
    """Returns the torch.device this rank drives (one GPU per process, or cpu when there is no GPU)."""
    import torch

    if torch.cuda.is_available():
        dev = device_index_for(info.local_rank)
        torch.cuda.set_device(dev)
        bind_cpu_near_gpu(dev)
        return torch.device("cuda", dev)
    return torch.device("cpu")


def bind_cpu_near_gpu(dev: int) -> bool:
    """Pins this process to the CPU cores NVML reports as local to its GPU, so pinned host batches
    are allocated on the GPU's NUMA node and H2D copies do not cross the socket interconnect
    (matters when 8 ranks feed 8 GPUs at once).  Best effort: silently skipped when NVML or the
    affinity call is unavailable."""
    try:
        import pynvml
        pynvml.nvmlInit()
        import torch
        uuid = torch.cuda.get_device_properties(dev).uuid
        h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + str(uuid)).encode())
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = {64 * i + b for i, w in enumerate(mask) for b in range(64) if (w >> b) & 1}
        allowed = os.sched_getaffinity(0)
        cpus &= allowed
        if cpus and cpus != allowed:
            os.sched_setaffinity(0, cpus)
            return True
    except Exception:
        pass
    return False


def init_process_group(info: RankInfo, backend: Optional[str] = None, timeout_s: int = 600):
    """Host-side rendezvous (bootstrap, object broadcast, IPC-handle exchange).  On a GPU box the
    group is NCCL (it is also the measured baseline for gradient reduction); on CPU it is gloo."""
    import datetime

    import torch
    import torch.distributed as dist

    if dist.is_initialized():
        return
    if backend is None:
        # one group, two transports: gloo for host objects, NCCL for device tensors
        backend = "cpu:gloo,cuda:nccl" if torch.cuda.is_available() else "gloo"
    os.environ.setdefault("MASTER_ADDR", info.master_addr)
    os.environ.setdefault("MASTER_PORT", str(info.master_port))
    kw = {}
    if "nccl" in backend:
        kw["device_id"] = torch.device("cuda", torch.cuda.current_device())
    dist.init_process_group(
        backend=backend, rank=info.rank, world_size=info.world_size,
        timeout=datetime.timedelta(seconds=timeout_s), **kw)
