"""Host-side pieces of b200dp.checkpoint (no GPU): the MinSizePartitioner restatement
(imagenet-resnet50-ps.py:75-78), ShardedVariable's axis-0 split and part_<i> naming, and the
re-assembly of a partitioned .npz."""
import os
import tempfile

import numpy as np

import b200dp  # noqa: F401
from b200dp import checkpoint as ck


def test_min_size_partitioner_matches_the_published_rule():
    p = ck.MinSizePartitioner(min_shard_bytes=256 << 10, max_shards=2)      # the script's values, NUM_PS = 2
    assert p((64,)) == 1                                 # 256 B bias: one shard
    assert p((7, 7, 3, 64)) == 1                         # 37,632 B: below the minimum
    assert p((3, 3, 512, 512)) == 2                      # 9.4 MB: as many as allowed
    assert p((2048, 1000)) == 2
    assert p((1, 1, 64, 256)) == 1                       # axis 0 has one row: cannot split
    q = ck.MinSizePartitioner(min_shard_bytes=256 << 10, max_shards=64)
    # dense kernel 2048 x 1000 fp32: 4000 B per row -> 66 rows per shard -> ceil(2048/66) = 32 shards
    assert q((2048, 1000)) == 32
    assert q((3, 3, 512, 512)) == 3                      # only 3 rows on axis 0
    assert ck.MinSizePartitioner(1, 8)((5, 2)) == 5


def test_split_and_read_reassemble():
    a = np.arange(7 * 3, dtype=np.float32).reshape(7, 3)
    parts = ck.split_axis0(a, 3)
    assert [p.shape[0] for p in parts] == [3, 2, 2]      # first dim % n shards are one row longer
    with tempfile.TemporaryDirectory() as d:
        path = os.path.join(d, "c.npz")
        np.savez(path, **{"x/part_%d" % i: p for i, p in enumerate(parts)}, y=np.float32(3), **{"optimizer/iterations": np.int64(5)})
        back = ck.read(path)
    assert np.array_equal(back["x"], a) and back["y"] == 3 and int(back["optimizer/iterations"]) == 5
