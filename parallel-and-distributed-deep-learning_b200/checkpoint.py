"""Checkpoint of a plan whose optimizer state is sharded across ranks.

Every reference script ends with ``model.save('ImageNet-...-reuse.h5')``
(imagenet-resnet50-mirror.py:83-86; rank 0 only under Horovod, imagenet-resnet50-hvd.py:125-129),
and the parameter-server script shards its variables with
``MinSizePartitioner(min_shard_bytes=256 << 10, max_shards=NUM_PS)`` (imagenet-resnet50-ps.py:75-78).
With b200dp the weights are mirrored but Adam's m / v (or the SGD accumulator) live as 1/n shards,
one per rank, so saving means: gather the shards (b2dp_gather_state), split the flat buffers per
variable (b2dp_unpack), name the pieces, write a file.  h5py is not installed on this image, so
the container is NumPy's ``.npz``; the names are the Keras variable names:

    <var>                         weights, e.g. conv1_conv/kernel            shape as in the model
    optimizer/<var>/m, .../v      Adam slots (SGD: optimizer/<var>/momentum)
    optimizer/iterations          int64 step count
    <var>/part_<i>                with a partitioner: axis-0 pieces, ShardedVariable's naming

Element order inside a variable is the memory order of the tensor the caller handed to the plan
(HWIO under Keras; the torch stand-in's own order there), reshaped to the variable's shape.

``load`` puts a checkpoint back into a plan built for ANY number of ranks: weights are mirrored
again and each rank keeps its own slice of the slots, so a job saved at n=4 continues at n=2 or 1.

save and load are collective (every rank of the plan calls them); only `writer_rank` touches the
file.  They run between steps, on the plan's stream, and are not part of the hot path.
"""
import math

import numpy as np

from ._lib import BUF_REDUCED, BUF_STATE0, BUF_STATE1, BUF_WEIGHTS
from .resnet50_shapes import resnet50_trainable_variables


class MinSizePartitioner:
    """tf.distribute.experimental.partitioners.MinSizePartitioner [upstream
    tensorflow/python/distribute/sharded_variable.py, partitioned_variables.min_max_variable_partitioner]:
    as many axis-0 shards as possible, at most `max_shards`, each at least `min_shard_bytes`."""

    def __init__(self, min_shard_bytes=256 << 10, max_shards=1, bytes_per_string=16):
        if min_shard_bytes < 1:
            raise ValueError("Argument `min_shard_bytes` must be positive. Received: %r" % (min_shard_bytes,))
        if max_shards < 1:
            raise ValueError("Argument `max_shards` must be positive. Received: %r" % (max_shards,))
        self.min_shard_bytes, self.max_shards = min_shard_bytes, max_shards

    def __call__(self, shape, itemsize=4):
        """Number of shards along axis 0 for a variable of `shape`."""
        if not shape or shape[0] <= 1:
            return 1
        bytes_per_slice = math.prod(shape[1:]) * itemsize
        slices_per_shard = max(1, math.ceil(self.min_shard_bytes / bytes_per_slice))
        return max(1, min(self.max_shards, math.ceil(shape[0] / slices_per_shard)))


def split_axis0(arr, shards):
    """ShardedVariable's split: axis 0 cut as evenly as possible, the first dim%n shards one row longer."""
    return np.array_split(arr, shards, axis=0) if shards > 1 else [arr]


def _variables(plan, variables):
    if variables is None:
        variables = resnet50_trainable_variables()
    variables = [(n, tuple(int(d) for d in s)) for n, s in variables]
    if [math.prod(s) for _, s in variables] != plan.layout.numels:
        raise ValueError("checkpoint: the variable list does not match the plan's tensor sizes")
    return variables


def _flat_to_vars(plan, dev, which, variables):
    """unpack(which) into host arrays, one per variable."""
    import torch
    device = "cuda:%d" % plan.ctx.devices[dev]
    outs = [torch.empty(math.prod(s), device=device) for _, s in variables]
    with torch.cuda.device(device):
        plan.unpack(dev, outs, which=which)
        torch.cuda.synchronize()
    return [o.cpu().numpy().reshape(s) for o, (_, s) in zip(outs, variables)]


def save(plan, path, step, optimizer="adam", variables=None, partitioner=None, writer_rank=0):
    """Collective.  `step` = optimizer iterations done so far.  Local contexts: drives every rank
    itself.  Returns the dict written (on the writer) or None."""
    variables = _variables(plan, variables)
    devs = range(len(plan.ctx.ranks))
    slots = {"m": BUF_STATE0, "v": BUF_STATE1} if optimizer == "adam" else {"momentum": BUF_STATE0}
    writer_dev = plan.ctx.ranks.index(writer_rank) if writer_rank in plan.ctx.ranks else None
    out = {}
    if writer_dev is not None:
        for (name, _), a in zip(variables, _flat_to_vars(plan, writer_dev, BUF_WEIGHTS, variables)):
            out[name] = a
    import torch
    for slot, which in slots.items():
        for d in devs:                       # the gather is a collective: issue it for every local rank
            with torch.cuda.device(plan.ctx.devices[d]):
                plan.gather_state(d, which)
        if writer_dev is not None:
            for (name, _), a in zip(variables, _flat_to_vars(plan, writer_dev, BUF_REDUCED, variables)):
                out["optimizer/%s/%s" % (name, slot)] = a
        for d in devs:                       # nobody starts the next gather before the writer has read this one
            torch.cuda.synchronize(plan.ctx.devices[d])
    if writer_dev is None:
        return None
    out["optimizer/iterations"] = np.int64(step)
    if partitioner is not None:
        for key in list(out):
            a = out[key]
            shards = partitioner(a.shape, a.dtype.itemsize) if isinstance(a, np.ndarray) and a.ndim else 1
            if shards > 1:
                del out[key]
                for i, part in enumerate(split_axis0(a, shards)):
                    out["%s/part_%d" % (key, i)] = part
    np.savez(path, **out)
    return out


def read(path):
    """The checkpoint as {name: array} with partitioned variables re-assembled."""
    raw = dict(np.load(path))
    out, parts = {}, {}
    for key, a in raw.items():
        head, _, tail = key.rpartition("/")
        if tail.startswith("part_") and tail[5:].isdigit():
            parts.setdefault(head, {})[int(tail[5:])] = a
        else:
            out[key] = a
    for head, pieces in parts.items():
        if sorted(pieces) != list(range(len(pieces))):
            raise ValueError("checkpoint: variable %s is missing partitions (have %s)" % (head, sorted(pieces)))
        out[head] = np.concatenate([pieces[i] for i in range(len(pieces))], axis=0)
    return out


def load(plan, path, optimizer="adam", variables=None):
    """Collective over the plan's local ranks (each process reads the file itself).  Mirrors the
    weights into every local rank and hands each rank its slice of the optimizer slots, whatever
    rank count the checkpoint was written with.  Returns the saved iteration count."""
    import torch
    variables = _variables(plan, variables)
    ck = read(path)
    lay = plan.layout
    slots = {"m": BUF_STATE0, "v": BUF_STATE1} if optimizer == "adam" else {"momentum": BUF_STATE0}

    def flat_padded(slot):
        flat = np.zeros(lay.total_elems, np.float32)
        for k, (name, shape) in enumerate(variables):
            a = ck["optimizer/%s/%s" % (name, slot)]
            if tuple(a.shape) != shape:
                raise ValueError("checkpoint: slot %s of %s has shape %s, the model expects %s" % (slot, name, a.shape, shape))
            off = lay.tensor_offset(k)
            flat[off:off + a.size] = a.astype(np.float32, copy=False).ravel()
        return flat

    for name, shape in variables:
        if name not in ck or tuple(ck[name].shape) != shape:
            raise ValueError("checkpoint: variable %s missing or of shape %s, the model expects %s"
                             % (name, tuple(ck[name].shape) if name in ck else None, shape))
    for d, rank in enumerate(plan.ctx.ranks):
        device = "cuda:%d" % plan.ctx.devices[d]
        with torch.cuda.device(device):
            ws = [torch.from_numpy(np.ascontiguousarray(ck[name], dtype=np.float32).ravel()).to(device) for name, _ in variables]
            plan.load_weights(d, ws)
            for slot, which in slots.items():
                flat = flat_padded(slot)
                shard = plan.tensor(d, which)
                for b in range(lay.num_buckets):
                    info = lay.bucket_info(b)
                    lo = info["offset"] + rank * info["slice"]
                    shard[info["shard_off"]:info["shard_off"] + info["slice"]].copy_(
                        torch.from_numpy(flat[lo:lo + info["slice"]]).to(device))
            torch.cuda.synchronize()
    return int(ck["optimizer/iterations"])
