"""The gradient list the hot path moves: Keras ResNet-50 v1 + Dense(1000).

Every reference script builds the same model (imagenet-resnet50-mirror.py:66-75,
imagenet-resnet50.py:52-61, -hvd.py:87-94, -multiworkers.py:80-89, -ps.py:127-136):
``keras.applications.ResNet50(include_top=False, weights=None, pooling='avg')``
followed by ``Dense(1000, softmax)``.  Keras' ResNet50 [upstream
keras/applications/resnet.py] has a bias on every convolution and trainable
gamma/beta on every BatchNorm, so ``model.trainable_variables`` is 214 tensors,
25,583,592 elements (torchvision's bias-free ResNet-50 is a different list).
Kernels are HWIO (kh, kw, cin, cout); the order below is layer-creation order.
"""
from math import prod

STACKS = ((64, 3, 1), (128, 4, 2), (256, 6, 2), (512, 3, 2))  # (filters, blocks, stride of first block)


def resnet50_trainable_variables(num_classes=1000):
    """[(name, shape)] for the 214 trainable variables."""
    out = []

    def conv(name, kh, kw, cin, cout):
        out.append((name + "/kernel", (kh, kw, cin, cout)))
        out.append((name + "/bias", (cout,)))

    def bn(name, c):
        out.append((name + "/gamma", (c,)))
        out.append((name + "/beta", (c,)))

    conv("conv1_conv", 7, 7, 3, 64)
    bn("conv1_bn", 64)
    cin = 64
    for si, (f, blocks, _stride) in enumerate(STACKS, start=2):
        for b in range(1, blocks + 1):
            p = "conv%d_block%d" % (si, b)
            if b == 1:  # projection shortcut
                conv(p + "_0_conv", 1, 1, cin, 4 * f)
                bn(p + "_0_bn", 4 * f)
            conv(p + "_1_conv", 1, 1, cin, f)
            bn(p + "_1_bn", f)
            conv(p + "_2_conv", 3, 3, f, f)
            bn(p + "_2_bn", f)
            conv(p + "_3_conv", 1, 1, f, 4 * f)
            bn(p + "_3_bn", 4 * f)
            cin = 4 * f
    out.append(("dense/kernel", (cin, num_classes)))
    out.append(("dense/bias", (num_classes,)))
    return out


def resnet50_numels(num_classes=1000):
    return [prod(s) for _, s in resnet50_trainable_variables(num_classes)]


NUM_TENSORS = 214
NUM_ELEMENTS = 25_583_592
NUM_BYTES_F32 = 4 * NUM_ELEMENTS
