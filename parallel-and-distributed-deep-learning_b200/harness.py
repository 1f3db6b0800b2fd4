"""Stand-in conv graph + training-step harness (torch/cuDNN), used by bench.py and smoke tests.

The north star keeps the convolution forward/backward "in the reference's TF/cuDNN graph".
TensorFlow is not installed on this image, so the graph that *produces* the
214 gradients and *consumes* the updated weights is the same Keras model
(imagenet-resnet50-mirror.py:66-75) rebuilt on torch + cuDNN: library code, not product.
Everything between "gradients exist" and "weights updated" -- the hot path -- goes through
libb200dp.so.  Deviations from the scripts, all outside the hot path: RandomCrop(244,244)
on a 224 input is ill-defined and is skipped; RandomFlip is skipped; softmax + sparse
categorical cross-entropy is computed as cross_entropy on logits.
"""
import math

import torch
import torch.nn as nn
import torch.nn.functional as Fn

from . import api
from ._lib import BUF_WEIGHTS
from .resnet50_shapes import STACKS, resnet50_numels


class _Block(nn.Module):
    """keras.applications.resnet block1: stride sits on the first 1x1 conv; every conv has a bias."""

    def __init__(self, cin, f, stride, project):
        super().__init__()
        self.proj = None
        if project:
            self.proj = nn.Conv2d(cin, 4 * f, 1, stride=stride, bias=True)
            self.proj_bn = nn.BatchNorm2d(4 * f, eps=1.001e-5)
        self.c1 = nn.Conv2d(cin, f, 1, stride=stride, bias=True)
        self.b1 = nn.BatchNorm2d(f, eps=1.001e-5)
        self.c2 = nn.Conv2d(f, f, 3, padding=1, bias=True)
        self.b2 = nn.BatchNorm2d(f, eps=1.001e-5)
        self.c3 = nn.Conv2d(f, 4 * f, 1, bias=True)
        self.b3 = nn.BatchNorm2d(4 * f, eps=1.001e-5)

    def ordered(self):
        mods = [self.proj, self.proj_bn] if self.proj is not None else []
        return mods + [self.c1, self.b1, self.c2, self.b2, self.c3, self.b3]

    def forward(self, x):
        sc = self.proj_bn(self.proj(x)) if self.proj is not None else x
        y = Fn.relu(self.b1(self.c1(x)))
        y = Fn.relu(self.b2(self.c2(y)))
        y = self.b3(self.c3(y))
        return Fn.relu(sc + y)


class KerasResNet50(nn.Module):
    """Rescaling(1/255) -> ResNet50(include_top=False, pooling='avg') -> Dense(1000)."""

    def __init__(self, num_classes=1000):
        super().__init__()
        self.conv1 = nn.Conv2d(3, 64, 7, stride=2, padding=3, bias=True)
        self.bn1 = nn.BatchNorm2d(64, eps=1.001e-5)
        blocks = []
        cin = 64
        for f, count, stride in STACKS:
            for b in range(count):
                blocks.append(_Block(cin, f, stride if b == 0 else 1, project=(b == 0)))
                cin = 4 * f
        self.blocks = nn.ModuleList(blocks)
        self.dense = nn.Linear(cin, num_classes)
        self.reset_keras_init()
        # base_model(x, training=False): BatchNorm normalises with its moving statistics,
        # which stay at their initial (0, 1); gamma/beta still receive gradients.
        for m in self.modules():
            if isinstance(m, nn.BatchNorm2d):
                m.eval()

    def train(self, mode=True):
        super().train(mode)
        for m in self.modules():
            if isinstance(m, nn.BatchNorm2d):
                nn.Module.train(m, False)
        return self

    def reset_keras_init(self):
        g = torch.Generator().manual_seed(42)
        for m in self.modules():
            if isinstance(m, (nn.Conv2d, nn.Linear)):
                w = m.weight
                rf = w[0][0].numel() if w.dim() > 2 else 1
                fan_in, fan_out = w.shape[1] * rf, w.shape[0] * rf
                lim = math.sqrt(6.0 / (fan_in + fan_out))   # glorot_uniform
                with torch.no_grad():
                    w.copy_((torch.rand(w.shape, generator=g) * 2 - 1) * lim)
                    m.bias.zero_()
            elif isinstance(m, nn.BatchNorm2d):
                with torch.no_grad():
                    m.weight.fill_(1.0)
                    m.bias.zero_()

    def keras_order_params(self):
        """Parameters in model.trainable_variables order (kernel, bias / gamma, beta per layer)."""
        mods = [self.conv1, self.bn1]
        for b in self.blocks:
            mods += b.ordered()
        mods.append(self.dense)
        out = []
        for m in mods:
            out += [m.weight, m.bias]
        return out

    def forward(self, x):
        x = x * (1.0 / 255.0)
        x = Fn.relu(self.bn1(self.conv1(x)))
        x = Fn.max_pool2d(x, 3, 2, padding=1)
        for b in self.blocks:
            x = b(x)
        x = torch.flatten(Fn.adaptive_avg_pool2d(x, 1), 1)
        return self.dense(x)


class DataParallelStep:
    """One training step of one rank: forward/backward on the stand-in graph, then the hot
    path (pack -> fused reduce-scatter + optimizer + all-gather) through libb200dp.so.

    The model's parameters are made views of the plan's flat weight buffer, so the
    all-gathered weights are what the next forward reads (no unpack pass)."""

    def __init__(self, model, plan, dev=0, optimizer="adam", lr=1e-3, mean_over_ranks=False, overlap=False,
                 params=None):
        self.model, self.plan, self.dev = model, plan, dev
        self.params = list(params) if params is not None else model.keras_order_params()
        numels = [p.numel() for p in self.params]
        if numels != plan.layout.numels:
            raise ValueError("model parameters do not match the plan's tensor list")
        flat = plan.tensor(dev, BUF_WEIGHTS)
        plan.load_weights(dev, [p.data for p in self.params])
        torch.cuda.current_stream().synchronize()
        for k, p in enumerate(self.params):
            off = plan.layout.tensor_offset(k)
            seg = flat[off:off + p.numel()]
            if p.dim() == 4 and not p.data.is_contiguous() and p.data.is_contiguous(memory_format=torch.channels_last):
                # keep channels_last conv kernels channels_last (dense OHWI memory): cuDNN then needs
                # no per-step layout copies; autograd gives the gradient the same strides
                o, i, h, w = p.shape
                p.data = seg.view(o, h, w, i).permute(0, 3, 1, 2)
            else:
                p.data = seg.view(p.shape)
        self._checked_strides = False
        self.optimizer, self.lr, self.t = optimizer, lr, 0
        self.postscale = 1.0 / plan.ctx.nranks if mean_over_ranks else 1.0
        self.overlap = bool(overlap)
        if self.overlap:
            self._init_overlap()

    # -- bucket pipeline overlapped with backward -----------------------
    # Gradients appear in reverse layer order during backward.  As soon as the last gradient
    # of a bucket exists, that bucket's pack + fused step is issued on a side stream, so the
    # reduction of the late layers runs under the convolutions of the early ones (what Horovod's
    # background fusion cycle does for hvd.DistributedOptimizer, imagenet-resnet50-hvd.py:101).
    # Every rank walks the same autograd graph, so buckets complete -- and the collective calls
    # are issued -- in the same order everywhere.
    def _init_overlap(self):
        lay = self.plan.layout
        self._bucket_of = [lay.tensor_bucket(k) for k in range(len(self.params))]
        self._bucket_size = [0] * lay.num_buckets
        for b in self._bucket_of:
            self._bucket_size[b] += 1
        self._left = list(self._bucket_size)
        self._side = torch.cuda.Stream(device=self.plan.ctx.devices[self.dev])
        self._hyper = None
        for k, p in enumerate(self.params):
            p.register_post_accumulate_grad_hook(lambda _p, k=k: self._on_grad(k))

    def _on_grad(self, k):
        b = self._bucket_of[k]
        self._left[b] -= 1
        if self._left[b] == 0:
            self._issue_bucket(b)

    def _issue_bucket(self, b):
        main = torch.cuda.current_stream()
        ready = torch.cuda.Event()
        ready.record(main)                 # every gradient of the bucket is enqueued before this point
        self._side.wait_event(ready)
        ptrs = [0 if p.grad is None else p.grad.data_ptr() for p in self.params]
        self.plan.pack_step(self.dev, api.ptr_array(ptrs), self.optimizer, self._hyper, bucket=b, stream=self._side)

    def _make_hyper(self):
        if self.optimizer == "adam":
            return api.adam_hyper(lr=self.lr, step=self.t, postscale=self.postscale)
        return api.sgdm_hyper(lr=self.lr, postscale=self.postscale)

    def forward_backward(self, images, labels):
        for p in self.params:
            p.grad = None
        logits = self.model(images)
        # Keras divides the per-replica loss by the replica count so that SUM gives the
        # global-batch mean gradient
        loss = Fn.cross_entropy(logits, labels) / (self.plan.ctx.nranks if self.postscale == 1.0 else 1)
        loss.backward()
        return loss

    def hot_path(self):
        self.t += 1
        if not self._checked_strides:   # flat position k of w and of g must be the same element
            for p in self.params:
                sig = lambda t: [st for st, sz in zip(t.stride(), t.shape) if sz > 1]   # size-1 dims carry no layout
                if sig(p.grad) != sig(p):
                    raise RuntimeError("gradient strides %s differ from parameter strides %s" % (p.grad.stride(), p.stride()))
            self._checked_strides = True
        if self.optimizer == "adam":
            h = api.adam_hyper(lr=self.lr, step=self.t, postscale=self.postscale)
        else:
            h = api.sgdm_hyper(lr=self.lr, postscale=self.postscale)
        self.plan.pack_step(self.dev, [p.grad for p in self.params], self.optimizer, h)

    def __call__(self, images, labels):
        if not self.overlap:
            loss = self.forward_backward(images, labels)
            self.hot_path()
            return loss
        self.t += 1
        self._hyper = self._make_hyper()
        self._left = list(self._bucket_size)
        loss = self.forward_backward(images, labels)     # the hooks issue each bucket as it completes
        if any(self._left):
            raise RuntimeError("backward produced no gradient for some tensors of a bucket")
        torch.cuda.current_stream().wait_stream(self._side)   # the next forward reads the updated weights
        return loss


def check_keras_shape(model):
    assert [p.numel() for p in model.keras_order_params()] == resnet50_numels()
