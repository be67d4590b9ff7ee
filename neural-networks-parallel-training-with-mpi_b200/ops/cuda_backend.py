"""CUDA backend of the trainer: wires the sm_100a kernels of libb200dp into the step schedule.

Per step (ref.py:164-211) on one GPU:

  tiny nets (fp32, e.g. the reference's 2-3-1)       wide / deep nets (bf16)
  -------------------------------------------       ----------------------------------------------
  1 launch: tiny_mlp fwd+MSE+bwd -> grad bucket      fwd: tcgen05 GEMM per hidden layer (bias+ReLU
  1 launch: fused all-reduce+SGD (or local SGD)           fused), head GEMV fused with MSE + dY
                                                     bwd: per layer wgrad GEMM (writes bf16 grads
                                                          straight into the bucket), column-sum for
                                                          db, dgrad GEMM with fused ReLU mask
                                                     sync: as soon as a layer's bucket is complete
                                                          its fused all-reduce+SGD kernel is queued
                                                          on a second stream and overlaps the rest
                                                          of backward.

No PyTorch op runs inside the step; torch only owns memory and streams.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Callable, List, Optional

import torch

from ..models.mlp import FlatMLP, MLPSpec
from ..parallel.grad_sync import GradSync, weights_and_divisor
from ..parallel.symm_mem import SymmetricBuffer
from . import _lib, allreduce_sgd as ar, gemm, kernels as K

TINY_MAX_DIM = 64
TINY_MAX_PARAMS = 16384


def _align(n: int, a: int = 256) -> int:
    return (n + a - 1) // a * a


def is_tiny(spec: MLPSpec) -> bool:
    return max(spec.sizes) <= TINY_MAX_DIM and spec.flat_numel <= TINY_MAX_PARAMS


# =============================================================================== compute: tiny
class TinyCompute:
    """Whole net in one SIMT launch (``csrc/tiny_mlp.cu``)."""
    name = "cuda-tiny"

    def __init__(self, flat: FlatMLP, max_rows: int):
        self.flat = flat
        self.desc = K.mlp_desc(flat.spec)
        dev = flat.device
        self.scratch = K.scratch(K.tiny_mlp_scratch_bytes(flat.spec, max(max_rows, 1)), dev)
        self.loss = torch.zeros((), dtype=torch.float32, device=dev)
        self.bucket_hook: Optional[Callable[[int], None]] = None

    def prepare_batch(self, x, y):
        return x.to(torch.float32).contiguous(), y.to(torch.float32).contiguous()

    def loss_and_grads(self, x, y) -> torch.Tensor:
        if x.shape[0] == 0:
            self.flat.grads.zero_()
            self.loss.fill_(float("nan"))
        else:
            K.tiny_mlp_fwd_bwd(self.desc, self.flat.params, x, y, self.flat.grads, self.loss, self.scratch)
        if self.bucket_hook is not None:
            for l in reversed(range(self.flat.spec.num_layers)):
                self.bucket_hook(l)
        return self.loss

    def predict(self, x):
        from . import reference as ref_ops
        return ref_ops.mlp_forward(self.flat, x.to(torch.float32)).to(torch.float32)


# =============================================================================== compute: gemm
class GemmCompute:
    """Tensor-core path: tcgen05 GEMMs + fused head/MSE kernels.  bf16 training runs ``kind::f16`` on
    the bf16 shadow weights (fp32 master kept by the optimizer); fp32 training - the reference's own
    precision, ref.py:159 - runs ``kind::tf32`` straight on the fp32 parameters."""
    name = "cuda-tcgen05"

    def __init__(self, flat: FlatMLP, max_rows: int):
        spec = flat.spec
        for s in spec.sizes[:-1]:
            if s % 8:
                raise ValueError(f"tensor-core path needs physical widths that are multiples of 8 (got {s}); "
                                 "build the spec with MLPSpec(sizes, pad_to=8)")
        self.nout = spec.sizes[-1]                    # physical outputs (padded to 8 when 5-7, 9-15, ...)
        self.nout_logical = spec.logical_sizes[-1]
        if self.nout > 4 and self.nout % 8:
            raise ValueError("the GEMM head needs a multiple of 8 outputs: build the spec with "
                             "MLPSpec(sizes, pad_to=8, pad_last=True)")
        self.flat, self.spec, self.L = flat, spec, spec.num_layers
        dev = flat.device
        R = max(max_rows, 1)
        self.max_rows = R
        self.cdt = cdt = torch.bfloat16 if flat.shadow is not None else torch.float32
        # activations a[l] = input of layer l (a[0] is the batch itself); d[l] = dLoss/da[l]
        self.acts: List[Optional[torch.Tensor]] = [None] + [
            torch.empty(R, spec.sizes[l], dtype=cdt, device=dev) for l in range(1, self.L)]
        hid = max(spec.sizes[1:self.L], default=8)
        self.dbuf = [torch.empty(R * hid, dtype=cdt, device=dev) for _ in range(2)]
        self.loss = torch.zeros((), dtype=torch.float32, device=dev)
        self.head_is_gemv = self.nout <= 4
        self.yhat = torch.empty(R, self.nout, dtype=torch.float32, device=dev)
        if self.head_is_gemv:
            self.dy32 = torch.empty(R, self.nout, dtype=torch.float32, device=dev)
            self.head_scratch = K.scratch(K.head_scratch_bytes(R, spec.sizes[self.L - 1], self.nout), dev)
        else:
            self.dyb = torch.empty(R, self.nout, dtype=cdt, device=dev)
            self.mse_scratch = K.scratch(K.mse_scratch_bytes(R * self.nout), dev)
            self.colsum_scratch = K.scratch(K.colsum_scratch_bytes(R, self.nout), dev)
        # fused bias gradients: the GEMM that produces dY_l also reduces its columns (tickets + partials)
        widest = max(spec.sizes[1:])
        self.gemm_cs_scratch = K.scratch(gemm.colsum_scratch_bytes(R, widest), dev)
        self.bucket_hook: Optional[Callable[..., None]] = None
        self.wgrad_chunks = {}            # layer -> [(row0, row1), ...] of its weight (set with the buckets)
        self.gemm_max_ctas = 0
        self.grad_scale = 1.0             # this rank's weight in the all-reduce, folded into dLoss/dYhat
        self._zero_targets = None
        if cdt == torch.float32:
            # kind::tf32 reads K-major operands only (an MN-major tf32 MMA is a no-op on B200), so the
            # fp32 path materialises backward's transposes: W^T for dgrad, dY^T and X^T for wgrad
            self.ldT = _align(R, 4)
            self.xT = torch.empty(max(spec.sizes[:-1]) * self.ldT, dtype=cdt, device=dev)
            self.dyT = torch.empty(max(spec.sizes[1:]) * self.ldT, dtype=cdt, device=dev)
            self.wT = torch.empty(max([spec.sizes[l] * spec.sizes[l + 1] for l in range(1, self.L)], default=16),
                                  dtype=cdt, device=dev)

    def prepare_batch(self, x, y):
        k_phys = self.spec.sizes[0]
        y = y.to(torch.float32)
        if y.shape[1] != self.nout:                    # padded head: zero targets for the padding outputs
            yp = torch.zeros(y.shape[0], self.nout, dtype=torch.float32, device=y.device)
            yp[:, :y.shape[1]].copy_(y)
            y = yp
        if x.shape[1] != k_phys:                       # padded input width: zero columns on the right
            xb = torch.zeros(x.shape[0], k_phys, dtype=self.cdt, device=x.device)
            xb[:, :x.shape[1]].copy_(x)
            return xb, y.contiguous()
        if self.cdt == torch.float32:
            return x.to(torch.float32).contiguous(), y.contiguous()
        xb = torch.empty(x.shape, dtype=torch.bfloat16, device=x.device)
        if x.numel():
            if x.dtype == torch.bfloat16:
                xb.copy_(x)
            else:
                K.cast_f32_bf16(x.to(torch.float32).contiguous(), xb)
        return xb, y.contiguous()

    # ---- views ----
    def _w(self, l):
        return self.flat.compute_weight(l)

    def _gw(self, l):
        return self.flat.weight(l, self.flat.grads)

    def _gb(self, l):
        return self.flat.bias(l, self.flat.grads)

    def _d(self, idx, rows, width):
        return self.dbuf[idx & 1][:rows * width].view(rows, width)

    def forward_hidden(self, x) -> torch.Tensor:
        rows = x.shape[0]
        h = x
        for l in range(self.L - 1):
            out = self.acts[l + 1][:rows]
            gemm.linear_fwd(h, self._w(l), self.flat.bias(l), relu=True, out=out, max_ctas=self.gemm_max_ctas)
            h = out
        return h

    def _dgrad(self, dy, l, a_in, out) -> None:
        """dLoss/d(input of layer l) with the ReLU mask of that input, and - fused into the epilogue -
        its column sums = the bias gradient of layer l-1 (the `sum` of AddmmBackward0, ref.py:176)."""
        kw = dict(out=out, max_ctas=self.gemm_max_ctas, colsum_out=self._gb(l - 1),
                  colsum_scratch=self.gemm_cs_scratch)
        w = self._w(l)
        if self.cdt == torch.float32:
            wT = self.wT[:w.numel()].view(w.shape[1], w.shape[0])
            K.transpose_f32(w, wT)
            gemm.gemm(dy, wT, out, mask_src=a_in, **{k: v for k, v in kw.items() if k != "out"})
        else:
            gemm.linear_dgrad(dy, w, a_in, **kw)

    def loss_and_grads(self, x, y) -> torch.Tensor:
        rows = x.shape[0]
        if rows == 0:
            self.flat.grads.zero_()
            self.loss.fill_(float("nan"))
            if self.bucket_hook is not None:
                for l in reversed(range(self.L)):
                    self.bucket_hook(l)
            return self.loss
        L, flat = self.L, self.flat
        h = self.forward_hidden(x)
        last = L - 1
        # ---- head forward + loss + dY, head backward ----
        dnext = None
        if self.head_is_gemv:
            yhat, dy = self.yhat[:rows], self.dy32[:rows]
            K.head_fwd_mse(h, self._w(last), flat.bias(last), y, yhat, dy, self.loss, self.head_scratch,
                           grad_scale=self.grad_scale)
            if last > 0:
                dnext = self._d(last, rows, self.spec.sizes[last])
            K.head_bwd(dy, self._w(last), h, dnext, self._gw(last), self._gb(last), self.head_scratch,
                       db_prev=self._gb(last - 1) if last > 0 else None)
            if self.bucket_hook is not None:
                self.bucket_hook(last)
        else:
            yhat, dy = self.yhat[:rows], self.dyb[:rows]
            gemm.linear_fwd(h, self._w(last), flat.bias(last), relu=False, out=yhat, max_ctas=self.gemm_max_ctas)
            # padded outputs predict 0 against a 0 target: they add nothing, the mean is over real outputs
            K.mse_loss_grad(yhat, y, dy, self.loss, self.mse_scratch, grad_scale=self.grad_scale,
                            denom=rows * self.nout_logical)
            K.colsum(dy, self._gb(last), self.colsum_scratch)
            if last > 0:
                dnext = self._d(last, rows, self.spec.sizes[last])
                self._dgrad(dy, last, h, dnext)
            self._wgrad_chunked(dy, h, last)
        # ---- hidden layers, top down ----
        for l in range(L - 2, -1, -1):
            dy_l = dnext                                   # dLoss/d(pre-activation of layer l), [rows, sizes[l+1]]
            a_in = x if l == 0 else self.acts[l][:rows]
            if l > 0:
                # dgrad first: it is on the critical path of backward, and it must have read W_l
                # before the first chunk of this layer's sync is allowed to update W_l
                dnext = self._d(l, rows, self.spec.sizes[l])
                self._dgrad(dy_l, l, a_in, dnext)
            self._wgrad_chunked(dy_l, a_in, l)
        return self.loss

    def _wgrad_chunked(self, dy, a_in, l) -> None:
        """dW_l = dY^T a_in, one GEMM per row chunk of W_l; each chunk's bucket is handed to the sync
        strategy as soon as its GEMM is queued.  (db_l was written by the kernel that produced dY.)"""
        gw = self._gw(l)
        chunks = self.wgrad_chunks.get(l) or [(0, gw.shape[0])]
        if self.cdt == torch.float32:
            rows = dy.shape[0]
            dyT = self.dyT[:dy.shape[1] * self.ldT].view(dy.shape[1], self.ldT)[:, :rows]
            xT = self.xT[:a_in.shape[1] * self.ldT].view(a_in.shape[1], self.ldT)[:, :rows]
            K.transpose_f32(dy, dyT)
            K.transpose_f32(a_in, xT)
        for c, (r0, r1) in enumerate(chunks):
            if self.cdt == torch.float32:
                gemm.gemm(dyT[r0:r1], xT, gw[r0:r1], max_ctas=self.gemm_max_ctas)
            elif len(chunks) == 1:
                gemm.linear_wgrad(dy, a_in, out=gw, max_ctas=self.gemm_max_ctas)
            else:
                gemm.linear_wgrad(dy[:, r0:r1], a_in, out=gw[r0:r1], max_ctas=self.gemm_max_ctas)
            if self.bucket_hook is not None:
                self.bucket_hook(l, c)

    @torch.no_grad()
    def predict(self, x) -> torch.Tensor:
        """Forward only, on the same kernels as training (the reference's dead eval stubs,
        ref.py:213-236)."""
        rows = x.shape[0]
        if rows > self.max_rows:
            return torch.cat([self.predict(x[i:i + self.max_rows]) for i in range(0, rows, self.max_rows)])
        if rows == 0:
            return torch.empty(0, self.nout_logical, dtype=torch.float32, device=x.device)
        if self._zero_targets is None:
            self._zero_targets = torch.zeros(self.max_rows, self.nout, dtype=torch.float32, device=x.device)
        xb, _ = self.prepare_batch(x, self._zero_targets[:rows])
        h = self.forward_hidden(xb)
        last = self.L - 1
        yhat = torch.empty(rows, self.nout, dtype=torch.float32, device=x.device)
        if self.head_is_gemv:
            loss = torch.empty((), dtype=torch.float32, device=x.device)
            K.head_fwd_mse(h, self._w(last), self.flat.bias(last), self._zero_targets[:rows], yhat,
                           self.dy32[:rows], loss, self.head_scratch)
        else:
            gemm.linear_fwd(h, self._w(last), self.flat.bias(last), relu=False, out=yhat,
                            max_ctas=self.gemm_max_ctas)
        return yhat[:, :self.nout_logical]


# =============================================================================== sync strategies
@dataclass
class Bucket:
    """A contiguous range of the flat buffers that is synchronised by one kernel.  ``layer`` /
    ``chunk`` say which piece of backward completes it: chunk ``c`` of layer ``l`` covers the weight
    rows ``[row0, row1)`` (the last chunk also carries the bias and the segment padding)."""
    layer: int
    chunk: int
    start: int
    stop: int
    row0: int = 0
    row1: int = 0


MIN_CHUNK_BLOCKS = 4     # >= 1024 weight rows per chunk: every chunk GEMM re-reads the whole activation
                         # matrix, so small chunks turn HBM-bound (measured: a 256-row chunk of the
                         # 4096x12288 wgrad ran at 0.43 PFLOP/s, profiles/timeline_wide_n1.md)


def _best_row_split(blocks: int, nb: int, units: int, kmax: int, min_blocks: int = MIN_CHUNK_BLOCKS) -> List[int]:
    """Splits ``blocks`` 256-row GEMM tile rows into <= kmax chunks (largest first, smallest last)
    such that the total number of GEMM waves over ``units`` CTA pairs does not grow, preferring the
    smallest last chunk (the part of the sync nothing can hide) and then more chunks."""
    import itertools

    def waves(b):
        return -(-(b * nb) // units)

    best = ([blocks], waves(blocks), blocks, 1)
    for k in range(2, max(2, kmax) + 1):
        if k > blocks:
            break
        for cuts in itertools.combinations(range(1, blocks), k - 1):
            parts = [b - a for a, b in zip((0,) + cuts, cuts + (blocks,))]
            if any(parts[i] < parts[i + 1] for i in range(k - 1)) or parts[-1] < min_blocks:
                continue                                     # non-increasing, no HBM-bound slivers
            w = sum(waves(p_) for p_ in parts)
            cand = (parts, w, parts[-1], k)
            if (w, cand[2], -k) < (best[1], best[2], -best[3]):
                best = cand
    return best[0]


def make_buckets(spec: MLPSpec, bucket_mb: float, elem_bytes: int, units: int = 74) -> List[Bucket]:
    """One bucket per layer; a layer whose gradient exceeds ``bucket_mb`` (default 32 MB) is cut into
    row chunks of its weight matrix: the wgrad GEMM runs chunk by chunk and each chunk's sync starts
    while the next chunk is still being computed.  That matters for the LAST layer of backward,
    whose sync has nothing else to hide behind.  Chunk sizes are chosen so the chunked GEMMs need
    no more waves over the ``units`` SM pairs than the unsplit GEMM.  Top layer first."""
    limit = int((bucket_mb if bucket_mb > 0 else 32.0) * (1 << 20))
    out: List[Bucket] = []
    for l in reversed(range(spec.num_layers)):
        s0, e0 = spec.layer_span(l)
        wseg, _ = spec.layer_segments(l)
        rows, cols = wseg.shape
        want = -(-(wseg.numel * elem_bytes) // limit)
        parts = [rows]
        if want > 1 and rows >= 512:
            blocks = -(-rows // 256)
            parts = [p_ * 256 for p_ in _best_row_split(blocks, -(-cols // 256), units, min(3, want))]
            parts[-1] -= sum(parts) - rows                      # the last chunk takes the ragged tail
            parts = [p_ for p_ in parts if p_ > 0]
        r = 0
        for c, n in enumerate(parts):
            r1 = r + n
            start = s0 + r * cols
            stop = e0 if r1 == rows else s0 + r1 * cols
            out.append(Bucket(l, c, start, stop, r, r1))
            r = r1
    return out


class BucketedSync(GradSync):
    """Common schedule of every GPU sync strategy: the step's gradient buffer is cut into buckets
    (one per layer by default); as soon as backward has produced a bucket (``bucket_ready``, called
    from inside the compute backend) its sync+update is queued on a side stream and overlaps the
    rest of backward; ``sync_and_step`` queues whatever is left and joins the streams."""

    def __init__(self, comm, weights, divisor, flat: FlatMLP, buckets: List[Bucket], overlap: bool):
        super().__init__(comm, weights, divisor)
        self.flat, self.buckets, self.overlap = flat, buckets, overlap
        self.comm_stream = torch.cuda.Stream(device=flat.device) if overlap else None
        self._ready = torch.cuda.Event()
        self._done = torch.cuda.Event()
        self._launched = set()
        self._lr = self._mu = None
        self._prof = None
        self._by_key = {(b.layer, b.chunk): i for i, b in enumerate(buckets)}
        self.chunks_of = {}
        for b in buckets:
            self.chunks_of.setdefault(b.layer, []).append(b)

    def _run_bucket(self, b: Bucket, lr: float, mu: float, stream) -> None:
        raise NotImplementedError

    # called by the compute backend right after chunk `c` of layer `l`'s backward has been queued
    # (c = None: everything of that layer)
    def bucket_ready(self, l: int, c: Optional[int] = None) -> None:
        if self._lr is None:
            return
        keys = [(l, c)] if c is not None else [(b.layer, b.chunk) for b in self.chunks_of.get(l, [])]
        for k in keys:
            i = self._by_key.get(k)
            if i is not None:
                self._launch(i)

    def _launch(self, i: int) -> None:
        if i in self._launched:
            return
        main = torch.cuda.current_stream()
        stream = main
        if self.comm_stream is not None:
            self._ready.record(main)
            self.comm_stream.wait_event(self._ready)
            stream = self.comm_stream
        if self._prof is not None:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            self._run_bucket(self.buckets[i], self._lr, self._mu, stream)
            e1.record(stream)
            self._prof[-1]["buckets"].append((e0, e1))
        else:
            self._run_bucket(self.buckets[i], self._lr, self._mu, stream)
        self._launched.add(i)

    # ---- "grad-sync us/step" (BASELINE.json metric) ----
    def start_profile(self) -> None:
        """From now on every step records CUDA events around each bucket's sync kernel and around the
        main stream's final wait for the side stream."""
        self._prof = []

    def stop_profile(self):
        """Per profiled step: ``busy_us`` = summed device time of the sync(+update) kernels,
        ``exposed_us`` = how long the main stream sat in the final join after its own last kernel
        (the part of the sync that backward did not hide)."""
        torch.cuda.synchronize()
        out = []
        for st in self._prof or []:
            busy = sum(a.elapsed_time(b) for a, b in st["buckets"]) * 1e3
            exposed = st["tail0"].elapsed_time(st["tail1"]) * 1e3 if "tail0" in st else busy
            out.append({"busy_us": busy, "exposed_us": exposed, "kernels": len(st["buckets"])})
        self._prof = None
        return out

    def begin_step(self, lr: float, mu: float) -> None:
        self._lr, self._mu = lr, mu
        self._launched.clear()
        if self._prof is not None:
            self._prof.append({"buckets": []})

    def sync_and_step(self, flat, lr, mu):
        if self._lr is None:
            self.begin_step(lr, mu)
        for i in range(len(self.buckets)):
            self._launch(i)
        if self.comm_stream is not None:
            main = torch.cuda.current_stream()
            self._done.record(self.comm_stream)
            if self._prof is not None:
                t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                t0.record(main)
                main.wait_event(self._done)
                t1.record(main)
                self._prof[-1]["tail0"], self._prof[-1]["tail1"] = t0, t1
            else:
                main.wait_event(self._done)
        self._lr = None
        flat.step_count += 1


class CudaLocalSync(BucketedSync):
    """World of one: the fused SGD kernel per bucket (HBM-bound; hidden under backward's GEMMs)."""
    name = "local"
    update_blocks = int(__import__("os").environ.get("B200DP_UPDATE_BLOCKS", "0"))

    def _run_bucket(self, b, lr, mu, stream):
        f = self.flat
        sl = slice(b.start, b.stop)
        K.sgd_momentum(f.params[sl], f.grads[sl], f.momentum[sl], f.shadow[sl] if f.shadow is not None else None,
                       lr, mu, first_step=False, grad_scale=1.0, stream=stream,
                       max_blocks=self.update_blocks if self.comm_stream is not None else 0)


class NcclAllReduceSync(BucketedSync):
    """Library baseline: ``torch.distributed.all_reduce`` (NCCL) per bucket, then the local SGD
    kernel.  Measured against, not shipped as the product path."""
    name = "allreduce"

    def _run_bucket(self, b, lr, mu, stream):
        import torch.distributed as dist
        f = self.flat
        sl = slice(b.start, b.stop)
        w = self.weights[self.comm.rank]
        with torch.cuda.stream(stream):
            g = f.grads[sl]
            if w != 1.0:
                g.mul_(w)
            dist.all_reduce(g, op=dist.ReduceOp.SUM)
        K.sgd_momentum(f.params[sl], f.grads[sl], f.momentum[sl], f.shadow[sl] if f.shadow is not None else None,
                       lr, mu, first_step=False, grad_scale=1.0 / self.divisor, stream=stream)


class CudaGatherRootSync(GradSync):
    """The reference's protocol (gather -> root average -> send, ref.py:185-203) with device
    tensors staged through the host communicator; baseline only."""
    name = "gather_root"

    def sync_and_step(self, flat, lr, mu):
        from . import reference as ref_ops
        comm = self.comm
        grad_list = comm.gather(flat.grads.detach().to("cpu"), root=0)
        if comm.rank == 0:
            avg = ref_ops.reduce_grads(grad_list, self.weights, self.divisor)
            for k in range(1, comm.size):
                comm.send(avg, k)
        else:
            avg = comm.recv(0)
        g = avg.to(flat.params.device, torch.float32)
        K.sgd_momentum(flat.params, g, flat.momentum, flat.shadow, lr, mu, first_step=False)
        flat.step_count += 1


class FusedSync(BucketedSync):
    """The product path: one fused all-reduce + SGD kernel per bucket over NVLink peer memory."""

    def __init__(self, comm, weights, divisor, mode: str, flat: FlatMLP, symm: "SymmLayout",
                 buckets: List[Bucket], overlap: bool, max_blocks: int, timeout_ms: int):
        super().__init__(comm, weights, divisor, flat, buckets, overlap)
        self.name = mode
        self.symm = symm
        dev = flat.device
        self.epochs = torch.zeros(ar.SYNC_MAX_BLOCKS, dtype=torch.int32, device=dev)
        self.error_flag = torch.zeros(1, dtype=torch.int32, device=dev)
        bf16 = flat.grads.dtype == torch.bfloat16
        m = ar.MODES[mode]
        # bf16 training ships only the bf16 shadow to the peers; the fp32 master of a slice lives
        # on its owner (gathered on demand).  fp32 training ships the fp32 values themselves.
        self.push_master = flat.shadow is None
        self.args = ar.make_args(
            comm.rank, comm.size, symm.ptrs("grads"), symm.ptrs("params"),
            symm.ptrs("shadow") if flat.shadow is not None else None, symm.ptrs("flags"),
            flat.momentum.data_ptr(), self.epochs.data_ptr(), self.error_flag.data_ptr(), weights, divisor,
            bf16, m, self.push_master, max_blocks, timeout_ms,
            symm.mc("grads"), symm.mc("shadow") if flat.shadow is not None else None, symm.mc("params"))
        self.sharded_master = (m != 0) and not self.push_master

        # kernels read the biases in fp32 from the master buffer: when the master is sharded, the part
        # of a bucket from its layer's bias onwards is still written into every rank
        self._tail = {}
        for i, b in enumerate(buckets):
            bias_off = flat.spec.layer_segments(b.layer)[1].offset
            self._tail[(b.start, b.stop)] = max(0, b.stop - max(b.start, bias_off)) if self.sharded_master else 0

    def _run_bucket(self, b, lr, mu, stream):
        ar.launch(self.args, b.start, b.stop - b.start, lr, mu, False, stream=stream,
                  replicate_tail=self._tail[(b.start, b.stop)])

    def calibrate(self, modes, iters: int = 4) -> str:
        """Start-up pick between sync variants: times each on the largest bucket (device-timed, max
        over ranks) and keeps the fastest - topology quirks (a 4-GPU subset of an 8-GPU NVSwitch
        domain, a box without NVLS) then cannot leave a slow default in place.  Runs with lr = 0,
        mu = 1 on the still-zero gradient buffer, so parameters and momentum are left bit-identical.
        Every rank executes the same launch sequence and adopts the same answer."""
        b = max(self.buckets, key=lambda k: k.stop - k.start)
        n = b.stop - b.start
        best, best_ms = self.name, float("inf")
        torch.cuda.synchronize()
        self.comm.barrier()                 # no rank may start spinning long before its peers arrive
        for mode in modes:
            self.args.mode = ar.MODES[mode]
            for _ in range(2):
                ar.launch(self.args, b.start, n, 0.0, 1.0, False)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            self.comm.barrier()
            e0.record()
            for _ in range(iters):
                ar.launch(self.args, b.start, n, 0.0, 1.0, False)
            e1.record()
            torch.cuda.synchronize()
            ms = max(self.comm.allgather(e0.elapsed_time(e1) / iters))
            if ms < best_ms:
                best, best_ms = mode, ms
        self.args.mode = ar.MODES[best]
        self.name = best
        self.check_errors()
        return best

    def check_errors(self) -> None:
        v = int(self.error_flag.item())
        if v:
            raise RuntimeError(f"fused gradient sync timed out waiting for rank {v - 1} "
                               "(peer dead or launch sequences diverged)")


class SymmLayout:
    """Carves one SymmetricBuffer into the regions the fused kernels need."""

    def __init__(self, comm, device, numel: int, grads_bf16: bool, with_shadow: bool, multicast: bool):
        self.regions = {}
        off = 0
        for name, nbytes in (("flags", ar.FLAG_BYTES), ("params", numel * 4),
                             ("grads", numel * (2 if grads_bf16 else 4)),
                             ("shadow", numel * 2 if with_shadow else 0)):
            self.regions[name] = (off, nbytes)
            off = _align(off + nbytes, 1024)
        self.buf = SymmetricBuffer(comm, off, device, flavour="auto", multicast=multicast)

    def ptrs(self, name: str) -> List[int]:
        off, _ = self.regions[name]
        return [p + off for p in self.buf.ptrs]

    def mc(self, name: str) -> Optional[int]:
        if self.buf.mc_ptr is None:
            return None
        return self.buf.mc_ptr + self.regions[name][0]

    def tensor(self, name: str, dtype) -> torch.Tensor:
        off, nbytes = self.regions[name]
        item = torch.empty((), dtype=dtype).element_size()
        return self.buf.tensor(dtype, nbytes // item, off)


# =============================================================================== runtime
class CudaRuntime:
    def __init__(self, trainer):
        cfg, comm, dev = trainer.cfg, trainer.comm, trainer.device
        _lib.lib()                                     # fail loudly if the native library is missing
        spec: MLPSpec = trainer.spec
        self.comm, self.cfg = comm, cfg
        tiny = is_tiny(spec) and cfg.dtype == "fp32"
        if not tiny:
            # tensor-core path: rows must be whole 16-byte vectors -> pad widths to multiples of 8; a
            # head with 5-7, 9-15, ... outputs is padded too (<= 4 outputs take the fused GEMV head)
            nout = spec.logical_sizes[-1]
            pad_last = nout > 4 and nout % 8 != 0
            if pad_last or any(s_ % 8 for s_ in spec.sizes[:-1]):
                spec = MLPSpec(spec.logical_sizes, pad_to=8, pad_last=pad_last)
                trainer.spec = spec
        sync_name = cfg.sync
        auto = sync_name == "auto"
        if auto:
            # measured on 8xB200 (profiles/allreduce_sweep_*): one-shot wins below ~256 KB, the
            # NVSwitch reduction (NVLS) wins at 8 GPUs (834 vs 628 GB/s), two-shot peer loads win at
            # 2 GPUs and tie at 4
            sync_name = "local"
            if comm.size > 1:
                sync_name = "fused_nvls" if comm.size > 4 else "fused_twoshot"
                if spec.flat_numel * 4 <= 256 * 1024:
                    sync_name = "fused_oneshot"
        fused = comm.size > 1 and sync_name.startswith("fused")
        weights, divisor = weights_and_divisor(trainer.counts, cfg.grad_weighting)
        uniform = all(w == weights[0] for w in weights)
        # Non-uniform per-rank weights (sample weighting of uneven shards): on the tensor-core
        # path every rank folds its weight into dLoss/dYhat, so its gradients arrive pre-weighted and any
        # reduction - the NVSwitch's plain sum included - gives sum_r w_r g_r.  The SIMT path keeps the
        # weights inside the peer-load kernels (one-/two-shot); it never takes NVLS with unequal weights.
        prescale = fused and not tiny and not uniform
        my_weight = weights[comm.rank]
        if prescale:
            weights = [1.0] * comm.size
        if fused and tiny and not uniform and sync_name == "fused_nvls":
            if comm.rank == 0 and not auto:
                print("b200dp: NVLS cannot weight its inputs - using fused_twoshot for the sample-weighted sync")
            sync_name = "fused_twoshot"
        self.symm: Optional[SymmLayout] = None
        grads_bf16 = cfg.dtype == "bf16"
        if fused:
            calibrate = auto and cfg.calibrate_sync and spec.flat_numel * 4 > 256 * 1024
            self.symm = SymmLayout(comm, dev, spec.flat_numel, grads_bf16, cfg.dtype == "bf16",
                                   multicast=(sync_name == "fused_nvls" or calibrate))
            if sync_name == "fused_nvls" and self.symm.buf.mc_ptr is None:
                if comm.rank == 0 and not auto:
                    print("b200dp: NVLS multicast is not available on this box - using fused_twoshot")
                sync_name = "fused_twoshot"
            self.flat = FlatMLP(spec, dev, cfg.dtype,
                                grads=self.symm.tensor("grads", torch.bfloat16 if grads_bf16 else torch.float32),
                                shadow=self.symm.tensor("shadow", torch.bfloat16) if cfg.dtype == "bf16" else None,
                                params=self.symm.tensor("params", torch.float32))
        else:
            self.flat = FlatMLP(spec, dev, cfg.dtype)
        max_rows = trainer.shard.rows
        self.compute = TinyCompute(self.flat, max_rows) if tiny else GemmCompute(self.flat, max_rows)
        if prescale:
            self.compute.grad_scale = float(my_weight)
        self.buckets = make_buckets(spec, cfg.bucket_mb, 2 if grads_bf16 else 4,
                                    units=max(torch.cuda.get_device_properties(dev).multi_processor_count // 2, 1))
        if tiny:
            self.buckets = [Bucket(0, 0, 0, spec.flat_numel)]
        self.device_broadcast = None
        overlap = cfg.overlap and not tiny
        if comm.size == 1:
            self.sync: GradSync = CudaLocalSync(comm, weights, divisor, self.flat, self.buckets, overlap)
        elif fused:
            self.sync = FusedSync(comm, weights, divisor, sync_name, self.flat, self.symm, self.buckets,
                                  overlap=overlap, max_blocks=int(__import__("os").environ.get("B200DP_SYNC_BLOCKS", "0")),
                                  timeout_ms=cfg.watchdog_ms)
            self.device_broadcast = self._device_broadcast
            if calibrate and self.symm.buf.mc_ptr is not None:
                sync_name = self.sync.calibrate(("fused_twoshot", "fused_nvls"))
        elif sync_name == "allreduce":
            self.sync = NcclAllReduceSync(comm, weights, divisor, self.flat, self.buckets, overlap)
        elif sync_name == "gather_root":
            self.sync = CudaGatherRootSync(comm, weights, divisor)
        else:
            raise ValueError(f"unknown sync mode {sync_name!r}")
        if isinstance(self.sync, BucketedSync):
            self.compute.bucket_hook = self.sync.bucket_ready
            if not tiny:
                self.compute.wgrad_chunks = {l: [(b.row0, b.row1) for b in bs]
                                             for l, bs in self.sync.chunks_of.items()}
        self.sync_name = sync_name

    def _device_broadcast(self, params: torch.Tensor) -> None:
        """K9: every rank pulls rank 0's parameters from peer memory (ref.py:87-88)."""
        torch.cuda.synchronize()
        self.comm.barrier()
        ar.broadcast(self.sync.args, 0, params.numel() * 4)
        torch.cuda.synchronize()
        self.comm.barrier()

    def gather_optimizer_state(self) -> None:
        """bf16 two-shot / NVLS keep the fp32 master and the momentum of a slice only on its owner
        (ZeRO-1 style).  Before a checkpoint the slices are collected on every rank."""
        sync = self.sync
        if not isinstance(sync, FusedSync):
            return
        import torch.distributed as dist
        torch.cuda.synchronize()
        world, rank = self.comm.size, self.comm.rank
        vn = 8 if self.flat.grads.dtype == torch.bfloat16 else 4
        if ar.MODES[sync.name] == 0:
            return
        for b in self.buckets:
            units = (b.stop - b.start) // vn
            for r in range(world):
                lo = b.start + units * r // world * vn
                hi = b.start + units * (r + 1) // world * vn
                if hi > lo:
                    dist.broadcast(self.flat.momentum[lo:hi], src=r)
                    if sync.sharded_master:
                        dist.broadcast(self.flat.params[lo:hi], src=r)
        torch.cuda.synchronize()

    def close(self) -> None:
        try:
            if isinstance(self.sync, FusedSync):
                self.sync.check_errors()
        finally:
            if self.symm is not None:
                self.symm.buf.close()
                self.symm = None


def build_cuda_runtime(trainer) -> CudaRuntime:
    return CudaRuntime(trainer)
