"""SPMD inference rules (N13) and the reshard function set (N12) of the semi-auto parallel API."""
import pytest

from paddle_b200.distributed.auto_parallel import spmd_rules as R
from paddle_b200.distributed.auto_parallel.spmd_rules import DistTensorSpec as S


def test_matmul_rules():
    # column parallel: x replicated, w split on n -> out split on n
    ins, (out,) = R.matmul(S((8, 16), [-1, -1]), S((16, 32), [-1, 0]))
    assert out.dims_mapping == [-1, 0] and not out.partial_dims
    # row parallel: contracted dim split on axis 0 -> partial output
    ins, (out,) = R.matmul(S((8, 16), [-1, 0]), S((16, 32), [0, -1]))
    assert out.dims_mapping == [-1, -1] and out.partial_dims == {0}
    # conflict on k (x says axis 0, w says axis 1): k is replicated, nothing partial
    ins, (out,) = R.matmul(S((8, 16), [-1, 0]), S((16, 32), [1, -1]))
    assert ins[0].dims_mapping == [-1, -1] and ins[1].dims_mapping == [-1, -1] and not out.partial_dims
    # data parallel batch + tensor parallel n, batched with broadcast, transposed weight
    ins, (out,) = R.matmul(S((4, 8, 16), [0, -1, -1]), S((32, 16), [1, -1]), trans_y=True)
    assert out.shape == (4, 8, 32) and out.dims_mapping == [0, -1, 1]
    # one mesh axis cannot shard two dims: m keeps axis 0, n loses it
    ins, (out,) = R.matmul(S((8, 16), [0, -1]), S((16, 32), [-1, 0]))
    assert out.dims_mapping == [0, -1] and ins[1].dims_mapping == [-1, -1]


def test_elementwise_reduction_embedding():
    ins, (out,) = R.elementwise(S((4, 8, 16), [0, -1, 1]), S((16,), [-1]))
    assert out.dims_mapping == [0, -1, 1] and ins[1].dims_mapping == [1]
    ins, (out,) = R.elementwise(S((4, 1, 16), [0, -1, -1]), S((4, 8, 16), [-1, 1, -1]))
    assert out.shape == (4, 8, 16) and out.dims_mapping == [0, 1, -1]
    ins, (out,) = R.reduction(S((4, 8, 16), [0, -1, 1]), axis=-1)
    assert out.shape == (4, 8) and out.dims_mapping == [0, -1] and out.partial_dims == {1}
    ins, (out,) = R.reduction(S((4, 8), [0, -1]), axis=0, keepdim=True)
    assert out.shape == (1, 8) and out.dims_mapping == [-1, -1] and out.partial_dims == {0}
    ins, (out,) = R.embedding(S((4, 8), [0, -1]), S((1000, 64), [1, -1]))          # vocab parallel
    assert out.shape == (4, 8, 64) and out.dims_mapping == [0, -1, -1] and out.partial_dims == {1}
    ins, (out,) = R.embedding(S((4, 8), [-1, -1]), S((1000, 64), [-1, 0]))         # hidden split
    assert out.dims_mapping == [-1, -1, 0] and not out.partial_dims


def test_norm_softmax_ce_layout_rules():
    ins, (y, mean, var) = R.layer_norm(S((4, 8, 16), [0, 1, 1]), S((16,), [-1]), S((16,), [-1]))
    assert ins[0].dims_mapping == [0, 1, -1] and y.dims_mapping == [0, 1, -1] and mean.shape == (4, 8)
    ins, (out,) = R.softmax(S((4, 8, 16), [0, -1, 1]), axis=-1)
    assert ins[0].dims_mapping == [0, -1, -1]
    ins, (sm, loss) = R.cross_entropy_with_softmax(S((32, 1000), [0, 1]), S((32, 1), [0, -1]))
    assert sm.dims_mapping == [0, 1] and loss.shape == (32, 1) and loss.dims_mapping == [0, -1]
    ins, (out,) = R.transpose(S((4, 8, 16), [0, -1, 1]), [2, 0, 1])
    assert out.shape == (16, 4, 8) and out.dims_mapping == [1, 0, -1]
    ins, (out,) = R.reshape(S((4, 8, 16), [0, -1, 1]), [32, 16])
    assert out.dims_mapping == [0, 1] and ins[0].dims_mapping == [0, -1, 1]
    ins, (out,) = R.reshape(S((4, 8, 16), [-1, 0, -1]), [32, 16])                    # a non-leading dim of a flatten
    assert ins[0].dims_mapping == [-1, -1, -1] and out.dims_mapping == [-1, -1]
    ins, (out,) = R.reshape(S((32, 16), [0, -1]), [4, 8, -1])
    assert out.shape == (4, 8, 16) and out.dims_mapping == [0, -1, -1]
    ins, outs = R.split(S((4, 8, 48), [0, -1, 1]), 3, axis=-1)
    assert ins[0].dims_mapping == [0, -1, -1] and len(outs) == 3 and outs[0].shape == (4, 8, 16)
    ins, (out,) = R.concat([S((4, 8), [0, -1]), S((4, 8), [-1, -1])], axis=1)
    assert out.shape == (4, 16) and out.dims_mapping == [0, -1] and ins[1].dims_mapping == [0, -1]
    ins, (out, lse) = R.flash_attention(S((2, 128, 8, 64), [0, -1, 1, -1]), S((2, 128, 8, 64), [0, 1, 1, -1]), S((2, 128, 8, 64), [0, -1, 1, -1]))
    assert out.dims_mapping == [0, -1, 1, -1] and ins[1].dims_mapping == [0, -1, 1, -1] and lse.dims_mapping == [0, 1, -1]
    ins, (out,) = R.get_spmd_rule("some_unknown_op")(S((4, 4), [0, 1]))
    assert out.dims_mapping == [-1, -1]


def test_spec_placement_round_trip():
    from paddle_b200.distributed.auto_parallel import Partial, Replicate, Shard
    sp = S((8, 16), [1, -1], {0})
    pl = sp.placements(2)
    assert isinstance(pl[0], Partial) and pl[1] == Shard(0)
    back = S.from_placements((8, 16), pl)
    assert back.dims_mapping == [1, -1] and back.partial_dims == {0}
    assert S.from_placements((8,), [Replicate()]).dims_mapping == [-1]
