"""Bucket / chunk planning (pure Python, no GPU)."""
import pytest

from b200dp.models.mlp import MLPSpec
from b200dp.ops.cuda_backend import _best_row_split, is_tiny, make_buckets


def _waves(parts, nb, units):
    return sum(-(-(p * nb) // units) for p in parts)


@pytest.mark.parametrize("blocks,nb", [(16, 48), (16, 16), (32, 32), (7, 5), (2, 100)])
def test_row_split_never_adds_gemm_waves(blocks, nb):
    parts = _best_row_split(blocks, nb, 74, 4)
    assert sum(parts) == blocks and all(p > 0 for p in parts)
    assert parts == sorted(parts, reverse=True)
    assert _waves(parts, nb, 74) <= _waves([blocks], nb, 74)


def test_buckets_cover_flat_buffer_exactly_once():
    for sizes in ([12288, 4096, 4096, 4096, 4096, 1], [4096] + [8192] * 4 + [1], [64, 256, 128, 8], [2, 3, 1]):
        spec = MLPSpec(sizes)
        bs = sorted(make_buckets(spec, 0, 2), key=lambda b: b.start)
        assert bs[0].start == 0 and bs[-1].stop == spec.flat_numel
        for a, b in zip(bs, bs[1:]):
            assert a.stop == b.start
        for b in bs:
            assert (b.stop - b.start) % 8 == 0 and b.start % 8 == 0
            w, _ = spec.layer_segments(b.layer)
            assert b.start == w.offset + b.row0 * w.shape[1]


def test_large_first_layer_is_chunked_small_layers_are_not():
    bs = make_buckets(MLPSpec([12288, 4096, 4096, 1]), 0, 2)
    per_layer = {}
    for b in bs:
        per_layer.setdefault(b.layer, []).append(b)
    assert len(per_layer[0]) > 1 and len(per_layer[1]) == 1 and len(per_layer[2]) == 1
    assert [b.layer for b in bs] == sorted([b.layer for b in bs], reverse=True)      # backward order
    last = per_layer[0][-1]
    assert 1024 <= last.row1 - last.row0 <= 1536                                      # small, but not HBM-bound, tail
    assert len(make_buckets(MLPSpec([12288, 4096, 1]), 1e6, 2)) == 2                 # limit disables chunking


def test_tiny_classification():
    assert is_tiny(MLPSpec([2, 3, 1])) and is_tiny(MLPSpec([7, 64, 64, 3]))
    assert not is_tiny(MLPSpec([128, 64, 1]))
