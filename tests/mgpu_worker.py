"""Worker for the multi-GPU tests: `torchrun --nproc-per-node N tests/mgpu_worker.py <scenario>`.
Prints one JSON line per rank."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

import b200dp  # noqa: E402,F401
from b200dp.config import TrainConfig  # noqa: E402
from b200dp.engine import Trainer  # noqa: E402
from b200dp.launch import bind_device, discover, init_process_group  # noqa: E402
from b200dp.parallel.comm import world_comm  # noqa: E402


class _Raw:
    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


def main():
    scenario = sys.argv[1]
    info = discover()
    dev = bind_device(info)
    init_process_group(info)
    comm = world_comm()
    out = {"rank": comm.rank, "scenario": scenario}
    if scenario.startswith("ref_"):
        sync = scenario[4:]
        t = Trainer(TrainConfig(quiet=True, sync=sync, nepochs=3, check_replicas_every=1), comm, dev)
        out["losses"] = [h["loss"] for h in t.train()]
        out["sync"] = t._runtime.sync_name
        out["fingerprint"] = t.flat.fingerprint()
        t.close()
    elif scenario.startswith("wide_"):
        sync = scenario[5:]
        cfg = TrainConfig(quiet=True, sync=sync, nepochs=6, n_samples=1024 * comm.size, n_features=256,
                          hidden=[512, 256], dtype="bf16", data_source="randn", data_dist="local", scale="none",
                          lr=1e-3, noise=0.1, init="torch_cpu")
        t = Trainer(cfg, comm, dev)
        out["losses"] = [h["loss"] for h in t.train()]
        out["sync"] = t._runtime.sync_name
        torch.cuda.synchronize()
        sh = t.flat.shadow.detach().cpu().view(torch.int16)
        import hashlib
        out["shadow_hash"] = hashlib.blake2b(sh.numpy().tobytes(), digest_size=8).hexdigest()
        ck = f"/tmp/b200dp_mgpu_{scenario}.pt"
        t.save_checkpoint(ck, 6)
        if comm.rank == 0:
            out["ckpt_norm"] = float(sum(v.double().norm() for v in torch.load(ck, weights_only=False)["model"].values()))
        t.close()
    elif scenario == "barrier":
        # K10: stream-ordered cross-GPU barrier.  Every rank publishes a value in its peer-mapped buffer,
        # barriers on the stream, then reads every peer's value with a plain device copy.
        from b200dp.ops import allreduce_sgd as ar
        from b200dp.ops.cuda_backend import SymmLayout
        symm = SymmLayout(comm, dev, 1024, False, False, multicast=False)
        mine = symm.tensor("params", torch.float32)
        epochs = torch.zeros(ar.SYNC_MAX_BLOCKS, dtype=torch.int32, device=dev)
        err = torch.zeros(1, dtype=torch.int32, device=dev)
        a = ar.make_args(comm.rank, comm.size, symm.ptrs("grads"), symm.ptrs("params"), None, symm.ptrs("flags"),
                         mine.data_ptr(), epochs.data_ptr(), err.data_ptr(), [1.0] * comm.size, 1.0, False, 0,
                         True, 1, 5000)
        seen = []
        for it in range(3):
            mine.fill_(float(100 * it + comm.rank))
            ar.peer_barrier(a)                       # everyone has written
            peer = (comm.rank + 1) % comm.size
            got = torch.empty(1, device=dev)
            raw = torch.as_tensor(_Raw(symm.ptrs("params")[peer], 4), device=dev).view(torch.float32)
            got.copy_(raw)
            ar.peer_barrier(a)                       # everyone has read before the next overwrite
            seen.append(float(got.item()))
        out["seen"] = seen
        out["err"] = int(err.item())
        symm.buf.close()
    elif scenario.startswith("uneven_"):
        sync = scenario[7:]
        t = Trainer(TrainConfig(quiet=True, sync=sync, nepochs=3, n_samples=17, grad_weighting="samples"), comm, dev)
        out["losses"] = [h["loss"] for h in t.train()]
        out["rows"] = t.shard.rows
        out["fingerprint"] = t.flat.fingerprint()
        t.close()
    elif scenario.startswith("oracle_"):
        # >= 1 MB of gradients (so auto would pick two-shot / NVLS), UNEVEN rows + sample weighting, 20
        # steps: bit-equal replicas (shadow AND fp32 biases) and parameters within 1e-3 of a single-
        # process fp32 model of ref.py:188-211.  lr 0.01: at 0.05 the fp32 model itself diverges here.
        import hashlib
        from b200dp.data import make_randn_rows
        from b200dp.models.mlp import MLP
        sync = scenario[7:]
        sizes, steps, lr, mu = [512, 2048, 512, 1], 20, 0.01, 0.9
        rows_total = 128 * comm.size + comm.size - 1
        cfg = TrainConfig(quiet=True, sync=sync, nepochs=steps, n_samples=rows_total, n_features=sizes[0],
                          hidden=sizes[1:-1], dtype="bf16", data_source="randn", data_dist="local", scale="none",
                          lr=lr, momentum=mu, noise=0.1, init="torch_cpu", grad_weighting="samples",
                          check_replicas_every=5)
        t = Trainer(cfg, comm, dev)
        out["losses"] = [h["loss"] for h in t.train()]
        out["sync"] = t._runtime.sync_name
        out["rows"] = t.shard.rows
        out["fingerprint"] = t.flat.fingerprint()
        hb = hashlib.blake2b(digest_size=8)
        for l in range(t.spec.num_layers):
            hb.update(t.flat.bias(l).detach().cpu().numpy().tobytes())
        out["bias_hash"] = hb.hexdigest()
        sd = t.state_dict()                       # collective: gathers the sharded fp32 master
        if comm.rank == 0:
            counts = list(t.counts)
            torch.manual_seed(cfg.seed)
            m = MLP(sizes).to(dev)
            m0 = {k: v.clone() for k, v in m.state_dict().items()}
            opt = torch.optim.SGD(m.parameters(), lr=lr, momentum=mu)
            params = list(m.parameters())
            shards, lo = [], 0
            for c in counts:
                d = make_randn_rows(lo, lo + c, sizes[0], 0.1, cfg.data_seed, 1)
                shards.append((torch.from_numpy(d[:, :sizes[0]]).float().to(dev),
                               torch.from_numpy(d[:, sizes[0]:]).float().to(dev)))
                lo += c
            for _ in range(steps):
                per_rank = []
                for x, y in shards:
                    opt.zero_grad()
                    torch.nn.functional.mse_loss(m(x), y).backward()
                    per_rank.append([p.grad.clone() for p in params])
                for k, p in enumerate(params):
                    acc = torch.zeros_like(p)
                    for r, g in enumerate(per_rank):
                        acc = acc + float(counts[r]) * g[k]
                    p.grad = acc / float(sum(counts))
                opt.step()
            out["err"] = {k: float((sd[k].to(dev) - v).abs().max()) for k, v in m.state_dict().items()}
            out["moved"] = {k: float((v - m0[k]).abs().max()) for k, v in m.state_dict().items()}
        t.close()
    sys.stdout.write("\nRESULT " + json.dumps(out) + "\n")     # one write: ranks share the pipe
    sys.stdout.flush()
    import torch.distributed as dist
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
