"""b200dp -- B200-native (sm_100a) data-parallel gradient reduction + optimizer step.

The one hot path of rrrickyz/Parallel-and-Distributed-Deep-Learning (per-step ResNet-50
gradient all-reduce + optimizer apply), behind the reference's own plug-in points.
Import name: ``b200dp`` (the root-level shim b200dp.py maps it to this directory).
"""
from ._lib import (BF16, BUF_GRADS, BUF_REDUCED, BUF_STATE0, BUF_STATE1, BUF_STATS, BUF_WEIGHTS, F32, OPT_ADAM,
                   OPT_NONE, OPT_SGDM, B200dpError, Hyper)
from .api import Context, Layout, Plan, adam_alpha, adam_apply_noalloc, adam_hyper, issue_group, ptr_array, sgdm_hyper
from .resnet50_shapes import (NUM_BYTES_F32, NUM_ELEMENTS, NUM_TENSORS, resnet50_numels,
                              resnet50_trainable_variables)

__all__ = [
    "BF16", "F32", "OPT_ADAM", "OPT_NONE", "OPT_SGDM", "BUF_GRADS", "BUF_WEIGHTS", "BUF_REDUCED", "BUF_STATE0",
    "BUF_STATE1", "BUF_STATS", "B200dpError", "Hyper", "Context", "Layout", "Plan", "adam_alpha", "adam_hyper",
    "sgdm_hyper", "ptr_array", "issue_group", "adam_apply_noalloc", "resnet50_numels", "resnet50_trainable_variables", "NUM_TENSORS", "NUM_ELEMENTS",
    "NUM_BYTES_F32",
]
