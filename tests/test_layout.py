"""The bucketing layer (b2dp_layout_*, host-only C) against a Python restatement of its
rule, on the ResNet-50 list and on ragged / empty / oversized edge cases."""
import numpy as np
import pytest

import b200dp

TENSOR_ALIGN, SLICE_ALIGN = 32, 256


def py_layout(numels, nranks, bucket_bytes):
    """Independent restatement: greedy in-order buckets by fp32 bytes, tensor starts on
    32 elements, bucket padded to nranks*256 elements."""
    up = lambda x, m: (x + m - 1) // m * m
    offsets, bucket_of, buckets = [], [], []
    cursor, cur = 0, None
    for k, n in enumerate(numels):
        if cur is not None and bucket_bytes > 0 and cur["count"] > 0 and cur["bytes"] + 4 * n > bucket_bytes:
            cur["elems"] = up(cursor - cur["offset"], nranks * SLICE_ALIGN)
            cursor = cur["offset"] + cur["elems"]
            buckets.append(cur)
            cur = None
        if cur is None:
            cur = {"first": k, "count": 0, "offset": cursor, "bytes": 0}
        cursor = up(cursor, TENSOR_ALIGN)
        offsets.append(cursor)
        bucket_of.append(len(buckets))
        cursor += n
        cur["count"] += 1
        cur["bytes"] += 4 * n
    if cur is not None:
        cur["elems"] = up(cursor - cur["offset"], nranks * SLICE_ALIGN)
        cursor = cur["offset"] + cur["elems"]
        buckets.append(cur)
    return offsets, bucket_of, buckets, cursor


CASES = [
    (b200dp.resnet50_numels(), 1, 0),
    (b200dp.resnet50_numels(), 8, 0),
    (b200dp.resnet50_numels(), 8, 25 << 20),
    (b200dp.resnet50_numels(), 4, 1 << 20),
    (b200dp.resnet50_numels(), 3, 64 << 20),
    ([64, 1000, 7, 0, 1027, 512], 4, 0),
    ([64, 1000, 7, 0, 1027, 512], 4, 4096),
    ([1], 8, 0),
    ([0, 0, 0], 2, 0),
    ([5, 3, 1 << 22, 9], 2, 1024),   # one tensor far larger than the bucket size
    ([], 2, 0),
]


@pytest.mark.parametrize("numels,nranks,bucket_bytes", CASES)
def test_layout_matches_restatement(numels, nranks, bucket_bytes):
    L = b200dp.Layout(numels, nranks=nranks, bucket_bytes=bucket_bytes)
    offsets, bucket_of, buckets, total = py_layout(numels, nranks, bucket_bytes)
    assert L.num_tensors == len(numels)
    assert L.num_buckets == len(buckets)
    assert L.total_elems == total
    assert [L.tensor_offset(k) for k in range(len(numels))] == offsets
    assert [L.tensor_bucket(k) for k in range(len(numels))] == bucket_of
    shard = 0
    for b, ref in enumerate(buckets):
        info = L.bucket_info(b)
        assert (info["first"], info["count"], info["offset"], info["elems"]) == (
            ref["first"], ref["count"], ref["offset"], ref["elems"])
        assert info["slice"] * nranks == info["elems"] and info["slice"] % SLICE_ALIGN == 0
        assert info["shard_off"] == shard
        shard += info["slice"]
    assert L.shard_elems == shard


def test_resnet50_layout_numbers():
    """214 tensors, 25,583,592 elements; N/8 is odd, so the flat buffer pads."""
    numels = b200dp.resnet50_numels()
    assert len(numels) == b200dp.NUM_TENSORS == 214 and sum(numels) == b200dp.NUM_ELEMENTS == 25_583_592
    L = b200dp.Layout(numels, nranks=8)
    assert L.num_buckets == 1
    assert L.total_elems >= sum(numels) and L.total_elems - sum(numels) < 8 * 256 + 214 * 32
    assert L.shard_elems * 8 == L.total_elems
    for k in range(214):
        assert L.tensor_offset(k) % 32 == 0


def test_no_tensor_overlap_and_all_inside_bucket():
    rng = np.random.default_rng(0)
    numels = [int(x) for x in rng.integers(0, 5000, 300)]
    L = b200dp.Layout(numels, nranks=5, bucket_bytes=100_000)
    end_prev = 0
    for k, n in enumerate(numels):
        off = L.tensor_offset(k)
        assert off >= end_prev
        end_prev = off + n
        info = L.bucket_info(L.tensor_bucket(k))
        assert info["offset"] <= off and off + n <= info["offset"] + info["elems"]


@pytest.mark.parametrize("kwargs", [dict(nranks=0), dict(nranks=9), dict(wire="fp8"), dict(bucket_bytes=-1)])
def test_bad_arguments_raise_valueerror(kwargs):
    with pytest.raises(ValueError):
        b200dp.Layout([16, 16], **kwargs)


def test_negative_numel_rejected():
    with pytest.raises(ValueError):
        b200dp.Layout([16, -1])


def test_out_of_range_queries():
    L = b200dp.Layout([16, 16])
    with pytest.raises(ValueError):
        L.bucket_info(1)
    with pytest.raises(b200dp.B200dpError):
        L.tensor_offset(2)


# ---------------------------------------------------------------- property-based (hypothesis)
from hypothesis import given, settings, strategies as st  # noqa: E402


@settings(max_examples=150, deadline=None)
@given(numels=st.lists(st.integers(min_value=0, max_value=20000), min_size=0, max_size=40),
       nranks=st.integers(min_value=1, max_value=8),
       bucket_bytes=st.sampled_from([0, 1, 1024, 4096, 65536, 1 << 20]),
       wire=st.sampled_from(["f32", "bf16"]))
def test_layout_invariants(numels, nranks, bucket_bytes, wire):
    """Whatever the tensor list: the C layout equals the restatement; tensors are ordered,
    128-byte aligned, disjoint and inside their bucket; buckets tile the flat range; every
    rank's slices are equal, 256-element multiples, and add up to the state shard."""
    L = b200dp.Layout(numels, nranks=nranks, wire=wire, bucket_bytes=bucket_bytes)
    offsets, bucket_of, buckets, total = py_layout(numels, nranks, bucket_bytes)
    assert L.total_elems == total and L.num_buckets == len(buckets)
    assert [L.tensor_offset(k) for k in range(len(numels))] == offsets
    cursor, shard, prev_end = 0, 0, 0
    for b in range(L.num_buckets):
        info = L.bucket_info(b)
        assert info["offset"] == cursor and info["elems"] % (nranks * 256) == 0
        assert info["slice"] * nranks == info["elems"] and info["shard_off"] == shard
        if bucket_bytes > 0 and info["count"] > 1:   # only a single oversized tensor may exceed the limit
            assert 4 * sum(numels[info["first"]:info["first"] + info["count"]]) <= bucket_bytes
        cursor += info["elems"]
        shard += info["slice"]
    assert cursor == L.total_elems and shard == L.shard_elems
    for k, n in enumerate(numels):
        off = L.tensor_offset(k)
        assert off % 32 == 0 and off >= prev_end
        prev_end = off + n
        info = L.bucket_info(L.tensor_bucket(k))
        assert info["offset"] <= off and off + n <= info["offset"] + info["elems"]
