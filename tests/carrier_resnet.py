"""Carrier network for BASELINE configs[2]: the reference's SimCLR ResNet-50 + MLP head, restated (test / bench
infrastructure only -- /root/reference does not exist on the GPU box, and the convolutions and linear layers are not
part of this library: they stay cuDNN / cuBLAS exactly as in the reference).

Follows /root/reference/networks/resnet_big.py: Bottleneck :36-66 (expansion 4, 1x1 -> 3x3(stride) -> 1x1, BN after
each conv, shortcut conv+BN when the shape changes, add -> ReLU), ResNet :69-118 (CIFAR stem: 3x3 stride-1 conv, no
max-pool; layers [3, 4, 6, 3] with 64/128/256/512 planes, strides 1/2/2/2; adaptive average pool; kaiming / constant
initialisation :84-89), SupConResNet :159-181 (head: Linear 2048 -> 2048, ReLU, Linear 2048 -> 128; the output is
L2-normalised, `F.normalize(self.head(feat), dim=1)`).

`fuse=True` builds the same network on this library's SyncBatchNorm with the activation folded into the batch-norm
kernels: `relu(bn(x))` becomes one pass (fuse_relu=True) and the block tail
`relu(bn3(conv3) + shortcut)` becomes `bn3(conv3, shortcut)` with the residual operand.  Parameter and buffer names
are identical either way, so state dicts are interchangeable.

`tail=True` additionally runs the end of the head through this library's ProjectionTail:
the last Linear, F.normalize and the stacking of the two views become one kernel and `forward` returns the
`[bsz, 2, feat_dim]` tensor the loss takes (the parameters stay `head.2.weight / head.2.bias`)."""
import torch
import torch.nn as nn
import torch.nn.functional as F


class Bottleneck(nn.Module):
    expansion = 4

    def __init__(self, in_planes, planes, stride=1, bn=nn.BatchNorm2d, fuse=False):
        super().__init__()
        kw = {'fuse_relu': True} if fuse else {}
        self.fuse = fuse
        self.conv1 = nn.Conv2d(in_planes, planes, kernel_size=1, bias=False)
        self.bn1 = bn(planes, **kw)
        self.conv2 = nn.Conv2d(planes, planes, kernel_size=3, stride=stride, padding=1, bias=False)
        self.bn2 = bn(planes, **kw)
        self.conv3 = nn.Conv2d(planes, self.expansion * planes, kernel_size=1, bias=False)
        self.bn3 = bn(self.expansion * planes, **kw)
        self.shortcut = nn.Sequential()
        if stride != 1 or in_planes != self.expansion * planes:
            self.shortcut = nn.Sequential(
                nn.Conv2d(in_planes, self.expansion * planes, kernel_size=1, stride=stride, bias=False),
                bn(self.expansion * planes))

    def forward(self, x):
        if self.fuse:
            out = self.bn1(self.conv1(x))                      # ReLU inside
            out = self.bn2(self.conv2(out))
            return self.bn3(self.conv3(out), self.shortcut(x))  # add + ReLU inside
        out = F.relu(self.bn1(self.conv1(x)))
        out = F.relu(self.bn2(self.conv2(out)))
        out = self.bn3(self.conv3(out))
        out = out + self.shortcut(x)
        return F.relu(out)


class SupConResNet50(nn.Module):
    def __init__(self, feat_dim=128, bn=nn.BatchNorm2d, fuse=False, width=64, blocks=(3, 4, 6, 3), tail=False):
        super().__init__()
        self.use_tail = tail
        kw = {'fuse_relu': True} if fuse else {}
        self.fuse = fuse
        self.in_planes = width
        self.conv1 = nn.Conv2d(3, width, kernel_size=3, stride=1, padding=1, bias=False)
        self.bn1 = bn(width, **kw)
        self.layer1 = self._make(width, blocks[0], 1, bn, fuse)
        self.layer2 = self._make(2 * width, blocks[1], 2, bn, fuse)
        self.layer3 = self._make(4 * width, blocks[2], 2, bn, fuse)
        self.layer4 = self._make(8 * width, blocks[3], 2, bn, fuse)
        dim_in = 8 * width * Bottleneck.expansion
        self.head = nn.Sequential(nn.Linear(dim_in, dim_in), nn.ReLU(inplace=True), nn.Linear(dim_in, feat_dim))
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode='fan_out', nonlinearity='relu')
            elif isinstance(m, nn.modules.batchnorm._BatchNorm):
                nn.init.constant_(m.weight, 1)
                nn.init.constant_(m.bias, 0)

    def _make(self, planes, n, stride, bn, fuse):
        layers = []
        for s in [stride] + [1] * (n - 1):
            layers.append(Bottleneck(self.in_planes, planes, s, bn, fuse))
            self.in_planes = planes * Bottleneck.expansion
        return nn.Sequential(*layers)

    def encode(self, x):
        out = self.bn1(self.conv1(x))
        if not self.fuse:
            out = F.relu(out)
        out = self.layer4(self.layer3(self.layer2(self.layer1(out))))
        return torch.flatten(F.adaptive_avg_pool2d(out, 1), 1)

    def forward(self, x, normalize=True):
        if self.use_tail:
            from simclr_pytorch_distributed_b200 import ProjectionTail
            tail = ProjectionTail.from_linear(self.head[2], n_views=2)      # shares head.2's parameters
            return tail(self.head[1](self.head[0](self.encode(x))))
        feat = self.head(self.encode(x))
        return F.normalize(feat, dim=1) if normalize else feat


def count_bn(model):
    return sum(1 for m in model.modules() if isinstance(m, nn.modules.batchnorm._BatchNorm))
