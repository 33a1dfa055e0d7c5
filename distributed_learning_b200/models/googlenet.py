"""GoogLeNet (Inception v1 with two auxiliary heads), defined here so nothing is downloaded.

Parity: reference ``src/network.py:33-54`` wraps ``torch.hub.load('pytorch/vision:v0.6.0',
'googlenet', pretrained=False, init_weights=False)``, which needs network access
(SURVEY.md Q15), and returns ``log_softmax(output[0])`` -- the main logits; the two
auxiliary heads are computed in training mode and discarded.  This module builds the same
architecture with the same parameter names, shapes and order as torchvision's
(187 tensors, 13 004 888 parameters; ``tests/test_core_cpu.py`` checks this against torchvision
when it is importable), and ``GoogLeNet.forward`` returns the same quantity.

The 14 tensors of ``aux1``/``aux2`` (6 379 984 parameters, 49 % of the gradient bytes) never
receive a gradient on this loss; the wrappers detect that at the first step and leave them out
of the fusion buckets instead of all-reducing uninitialised memory.
``compute_aux=False`` additionally skips their (dead) forward computation; the default keeps
it so the measured step does the same work as the reference's.
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F


class BasicConv2d(nn.Module):
    def __init__(self, cin: int, cout: int, **kw):
        super().__init__()
        self.conv = nn.Conv2d(cin, cout, bias=False, **kw)
        self.bn = nn.BatchNorm2d(cout, eps=0.001)

    def forward(self, x):
        return F.relu(self.bn(self.conv(x)), inplace=True)


class Inception(nn.Module):
    def __init__(self, cin, ch1x1, ch3x3red, ch3x3, ch5x5red, ch5x5, pool_proj):
        super().__init__()
        self.branch1 = BasicConv2d(cin, ch1x1, kernel_size=1)
        self.branch2 = nn.Sequential(BasicConv2d(cin, ch3x3red, kernel_size=1),
                                     BasicConv2d(ch3x3red, ch3x3, kernel_size=3, padding=1))
        # torchvision keeps the original implementation's 3x3 kernel in the "5x5" branch
        self.branch3 = nn.Sequential(BasicConv2d(cin, ch5x5red, kernel_size=1),
                                     BasicConv2d(ch5x5red, ch5x5, kernel_size=3, padding=1))
        self.branch4 = nn.Sequential(nn.MaxPool2d(kernel_size=3, stride=1, padding=1, ceil_mode=True),
                                     BasicConv2d(cin, pool_proj, kernel_size=1))

    def forward(self, x):
        return torch.cat([self.branch1(x), self.branch2(x), self.branch3(x), self.branch4(x)], 1)


class InceptionAux(nn.Module):
    def __init__(self, cin: int, num_classes: int, dropout: float = 0.7):
        super().__init__()
        self.conv = BasicConv2d(cin, 128, kernel_size=1)
        self.fc1 = nn.Linear(2048, 1024)
        self.fc2 = nn.Linear(1024, num_classes)
        self.dropout = nn.Dropout(p=dropout)

    def forward(self, x):
        x = F.adaptive_avg_pool2d(x, (4, 4))
        x = torch.flatten(self.conv(x), 1)
        x = self.dropout(F.relu(self.fc1(x), inplace=True))
        return self.fc2(x)


class GoogLeNetBackbone(nn.Module):
    """Returns ``(logits, aux2, aux1)`` in training mode with aux heads, else ``logits``."""

    def __init__(self, num_classes: int = 1000, aux_logits: bool = True, dropout: float = 0.2,
                 dropout_aux: float = 0.7):
        super().__init__()
        self.aux_logits = aux_logits
        self.conv1 = BasicConv2d(3, 64, kernel_size=7, stride=2, padding=3)
        self.maxpool1 = nn.MaxPool2d(3, stride=2, ceil_mode=True)
        self.conv2 = BasicConv2d(64, 64, kernel_size=1)
        self.conv3 = BasicConv2d(64, 192, kernel_size=3, padding=1)
        self.maxpool2 = nn.MaxPool2d(3, stride=2, ceil_mode=True)
        self.inception3a = Inception(192, 64, 96, 128, 16, 32, 32)
        self.inception3b = Inception(256, 128, 128, 192, 32, 96, 64)
        self.maxpool3 = nn.MaxPool2d(3, stride=2, ceil_mode=True)
        self.inception4a = Inception(480, 192, 96, 208, 16, 48, 64)
        self.inception4b = Inception(512, 160, 112, 224, 24, 64, 64)
        self.inception4c = Inception(512, 128, 128, 256, 24, 64, 64)
        self.inception4d = Inception(512, 112, 144, 288, 32, 64, 64)
        self.inception4e = Inception(528, 256, 160, 320, 32, 128, 128)
        self.maxpool4 = nn.MaxPool2d(2, stride=2, ceil_mode=True)
        self.inception5a = Inception(832, 256, 160, 320, 32, 128, 128)
        self.inception5b = Inception(832, 384, 192, 384, 48, 128, 128)
        if aux_logits:
            self.aux1 = InceptionAux(512, num_classes, dropout=dropout_aux)
            self.aux2 = InceptionAux(528, num_classes, dropout=dropout_aux)
        else:
            self.aux1 = None
            self.aux2 = None
        self.avgpool = nn.AdaptiveAvgPool2d((1, 1))
        self.dropout = nn.Dropout(p=dropout)
        self.fc = nn.Linear(1024, num_classes)

    def forward(self, x, compute_aux: bool = True):
        x = self.maxpool1(self.conv1(x))
        x = self.maxpool2(self.conv3(self.conv2(x)))
        x = self.inception3b(self.inception3a(x))
        x = self.maxpool3(x)
        x = self.inception4a(x)
        want_aux = self.aux_logits and self.training and compute_aux
        aux1 = self.aux1(x) if want_aux else None
        x = self.inception4d(self.inception4c(self.inception4b(x)))
        aux2 = self.aux2(x) if want_aux else None
        x = self.inception4e(x)
        x = self.maxpool4(x)
        x = self.inception5b(self.inception5a(x))
        x = torch.flatten(self.avgpool(x), 1)
        x = self.fc(self.dropout(x))
        if self.aux_logits and self.training:
            return x, aux2, aux1
        return x


class GoogLeNet(nn.Module):
    """The reference's model: log-softmax of the main logits of a GoogLeNet with aux heads."""

    def __init__(self, num_classes: int = 1000, compute_aux: bool = True):
        super().__init__()
        self.model = GoogLeNetBackbone(num_classes=num_classes, aux_logits=True)
        self.compute_aux = compute_aux

    def forward(self, x):
        out = self.model(x, compute_aux=self.compute_aux)
        logits = out[0] if isinstance(out, tuple) else out
        return F.log_softmax(logits, dim=1)
