"""The MLP regressor.

``MLP`` is the user-facing ``nn.Module`` with the reference's interface (``MLP()``, ``forward(x)``,
a ``layers`` ``nn.Sequential`` of Linear/ReLU pairs; ref.py:35-51) generalised to any width list.

``FlatMLP`` is what the trainer actually steps: one flat fp32 buffer for all parameters, one for
the gradients (the all-reduce bucket; peer-mapped on multi-GPU runs), one for the momentum and, for
bf16 training, a bf16 shadow copy of the weights that the GEMMs read.  Per-layer tensors are
views into those buffers.  Flat offsets follow ``model.parameters()`` order of the reference
(weight, bias per Linear; ref.py:181), each segment padded to 256 bytes so every weight matrix can
be described by a TMA tensor map and moved with 16-byte vectors.  ``state_dict()`` uses the
reference's key names (``layers.0.weight`` ...) so a checkpoint loads into the
reference's own ``MLP``.
"""
from __future__ import annotations

import math
from collections import OrderedDict
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence

import torch
from torch import nn

ALIGN_ELEMS = 128   # segment alignment in elements: 512 B in fp32, 256 B in bf16


class MLP(nn.Module):
    """Multilayer perceptron for regression; ``MLP()`` is the reference's 2-3-1 net (ref.py:41-45)."""

    def __init__(self, sizes: Sequence[int] = (2, 3, 1)):
        super().__init__()
        mods: List[nn.Module] = []
        for i in range(len(sizes) - 1):
            mods.append(nn.Linear(sizes[i], sizes[i + 1]))
            if i < len(sizes) - 2:
                mods.append(nn.ReLU())
        self.layers = nn.Sequential(*mods)
        self.sizes = tuple(sizes)

    def forward(self, x):
        return self.layers(x)


@dataclass(frozen=True)
class Segment:
    key: str
    layer: int
    kind: str        # "weight" | "bias"
    offset: int      # elements into the flat buffers
    numel: int
    shape: tuple     # physical shape inside the flat buffers (feature sizes rounded up to `pad_to`)
    logical_shape: tuple = ()   # the model's own shape (what state_dict exposes)


def _round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


class MLPSpec:
    """Layer sizes and the flat-buffer layout.  ``pad_to`` > 1 (the tensor-core path uses 8) rounds the
    input and hidden widths up so every row of every matrix is a whole number of 16-byte vectors;
    the padding rows / columns hold zeros, stay zero under training (their activations and
    gradients are identically zero) and never appear in ``state_dict``.  ``sizes`` are the physical
    (padded) widths the kernels see, ``logical_sizes`` the model's own."""

    def __init__(self, sizes: Sequence[int], pad_to: int = 1, pad_last: bool = False):
        if len(sizes) < 2:
            raise ValueError("an MLP needs at least an input and an output size")
        self.logical_sizes = tuple(int(s) for s in sizes)
        self.pad_to = int(pad_to)
        self.pad_last = bool(pad_last)     # also round the output width up (GEMM head with 5-7, 9-15, ... outputs)
        self.sizes = tuple(_round_up(s, self.pad_to) if (i < len(sizes) - 1 or pad_last) else s
                           for i, s in enumerate(self.logical_sizes))
        self.num_layers = len(self.sizes) - 1
        segs: List[Segment] = []
        off = 0
        for l in range(self.num_layers):
            fan_in, fan_out = self.sizes[l], self.sizes[l + 1]
            lin, lout = self.logical_sizes[l], self.logical_sizes[l + 1]
            for kind, shape, lshape in (("weight", (fan_out, fan_in), (lout, lin)), ("bias", (fan_out,), (lout,))):
                n = 1
                for s in shape:
                    n *= s
                segs.append(Segment(f"layers.{2 * l}.{kind}", l, kind, off, n, shape, lshape))
                off = _round_up(off + n, ALIGN_ELEMS)
        self.segments = tuple(segs)
        self.flat_numel = off
        self.num_params = sum(math.prod(s.logical_shape) for s in segs)
        self.padded = self.sizes != self.logical_sizes

    def layer_segments(self, l: int):
        return self.segments[2 * l], self.segments[2 * l + 1]

    def layer_span(self, l: int):
        """[start, stop) elements of layer ``l`` (weight + bias + padding): the per-layer bucket."""
        start = self.segments[2 * l].offset
        stop = self.segments[2 * l + 2].offset if l + 1 < self.num_layers else self.flat_numel
        return start, stop

    def is_tiny(self, limit: int = 64) -> bool:
        """True when every dimension is too small to feed a tensor-core tile (the reference's
        2-3-1 net): such models run on the fused SIMT kernel instead of the tcgen05 GEMMs."""
        return max(self.sizes) <= limit


class FlatMLP:
    def __init__(self, spec: MLPSpec, device, dtype: str = "fp32",
                 grads: Optional[torch.Tensor] = None, shadow: Optional[torch.Tensor] = None,
                 params: Optional[torch.Tensor] = None):
        self.spec, self.device, self.dtype = spec, torch.device(device), dtype
        n = spec.flat_numel
        self.params = params if params is not None else torch.zeros(n, dtype=torch.float32, device=device)
        self.momentum = torch.zeros(n, dtype=torch.float32, device=device)
        gdt = torch.bfloat16 if dtype == "bf16" else torch.float32
        self.grads = grads if grads is not None else torch.zeros(n, dtype=gdt, device=device)
        self.shadow = None
        if dtype == "bf16":
            self.shadow = shadow if shadow is not None else torch.zeros(n, dtype=torch.bfloat16, device=device)
        self.step_count = 0

    # ---- views ----
    def _view(self, flat: torch.Tensor, seg: Segment) -> torch.Tensor:
        return flat[seg.offset:seg.offset + seg.numel].view(seg.shape)

    def _logical(self, flat: torch.Tensor, seg: Segment) -> torch.Tensor:
        v = self._view(flat, seg)
        if seg.shape == seg.logical_shape:
            return v
        return v[tuple(slice(0, n) for n in seg.logical_shape)]

    def weight(self, l: int, flat: Optional[torch.Tensor] = None) -> torch.Tensor:
        return self._view(self.params if flat is None else flat, self.spec.segments[2 * l])

    def bias(self, l: int, flat: Optional[torch.Tensor] = None) -> torch.Tensor:
        return self._view(self.params if flat is None else flat, self.spec.segments[2 * l + 1])

    def compute_weight(self, l: int) -> torch.Tensor:
        """The tensor the GEMMs read: the bf16 shadow in bf16 mode, else the fp32 master."""
        return self.weight(l, self.shadow if self.shadow is not None else self.params)

    def compute_bias(self, l: int) -> torch.Tensor:
        return self.bias(l, self.params)

    # ---- initialisation ----
    def init_like_reference(self, seed: int = 0) -> None:
        """``torch.manual_seed(seed)`` then ``nn.Linear`` default init on the CPU, in module order -
        bit-identical to what root creates at ref.py:69,84 and broadcasts at ref.py:87."""
        gen_state = torch.random.get_rng_state()
        try:
            torch.manual_seed(seed)
            ref = MLP(self.spec.logical_sizes)
        finally:
            torch.random.set_rng_state(gen_state)
        self.load_state_dict(ref.state_dict())

    def init_on_device(self, seed: int = 0) -> None:
        """Same distribution (uniform +-1/sqrt(fan_in) for weight and bias), generated on the
        device: for billion-parameter models where a CPU init would dominate start-up."""
        g = torch.Generator(device=self.device).manual_seed(seed)
        self.params.zero_()                                  # padding rows / columns must be zero
        for seg in self.spec.segments:
            fan_in = self.spec.logical_sizes[seg.layer]
            bound = 1.0 / (fan_in ** 0.5)
            v = self._logical(self.params, seg)
            v.copy_((torch.rand(seg.logical_shape, generator=g, device=self.device) * 2 - 1) * bound)
        self.sync_shadow()

    def sync_shadow(self) -> None:
        if self.shadow is not None:
            self.shadow.copy_(self.params)

    # ---- state dict (reference key names) ----
    def state_dict(self) -> "OrderedDict[str, torch.Tensor]":
        out = OrderedDict()
        for seg in self.spec.segments:
            out[seg.key] = self._logical(self.params, seg).detach().to("cpu", copy=True).contiguous()
        return out

    def load_state_dict(self, sd: Dict[str, torch.Tensor]) -> None:
        keys = [s.key for s in self.spec.segments]
        missing = [k for k in keys if k not in sd]
        extra = [k for k in sd if k not in keys]
        if missing or extra:
            raise KeyError(f"state_dict mismatch: missing={missing} unexpected={extra}")
        for seg in self.spec.segments:
            t = sd[seg.key]
            if tuple(t.shape) != tuple(seg.logical_shape):
                raise ValueError(f"{seg.key}: shape {tuple(t.shape)} != {seg.logical_shape}")
            self._view(self.params, seg).zero_()
            self._logical(self.params, seg).copy_(t.to(self.device, torch.float32))
        self.sync_shadow()

    def momentum_state(self) -> "OrderedDict[str, torch.Tensor]":
        return OrderedDict((s.key, self._logical(self.momentum, s).detach().to("cpu", copy=True).contiguous())
                           for s in self.spec.segments)

    def load_momentum_state(self, sd: Dict[str, torch.Tensor]) -> None:
        for seg in self.spec.segments:
            self._view(self.momentum, seg).zero_()
            self._logical(self.momentum, seg).copy_(sd[seg.key].to(self.device, torch.float32))

    def as_module(self) -> MLP:
        m = MLP(self.spec.logical_sizes)
        m.load_state_dict(self.state_dict())
        return m

    def fingerprint(self) -> int:
        """Order-sensitive 64-bit hash of every byte a kernel computes with (replica-equality
        checks).  fp32 mode: the fp32 parameters.  bf16 mode: the bf16 shadow weights AND the fp32
        biases (the GEMM / head epilogues read those from the master buffer); the rest of the fp32
        master is left out because a slice of it may live only on its owner rank."""
        import hashlib

        h = hashlib.blake2b(digest_size=8)

        def feed(t: torch.Tensor) -> None:
            h.update(t.detach().to("cpu").contiguous().view(torch.uint8).numpy().tobytes())

        if self.shadow is None:
            feed(self.params)
        else:
            feed(self.shadow)
            for l in range(self.spec.num_layers):
                feed(self.bias(l))
        return int.from_bytes(h.digest(), "little")
