"""The small CNN of the reference, with the input shape as a parameter.

Parity: reference ``src/network.py:7-30`` (``BasicNet``): conv(C->32,3) relu conv(32->64,3)
relu maxpool2 dropout.25 flatten fc(->128) relu dropout.5 fc(128->10) log_softmax.  The
reference hard-wires 1x28x28 (``fc1`` input 9216 = 64*12*12); here ``in_channels`` and
``image_size`` are arguments so the CIFAR-10-shaped config is 3x32x32 with
``fc1`` input 64*14*14 = 12544.  With the defaults the parameter list is identical:
8 tensors, 1 199 882 parameters.
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F


class BasicNet(nn.Module):
    def __init__(self, in_channels: int = 1, image_size: int = 28, num_classes: int = 10):
        super().__init__()
        self.conv1 = nn.Conv2d(in_channels, 32, 3, 1)
        self.conv2 = nn.Conv2d(32, 64, 3, 1)
        self.dropout1 = nn.Dropout(0.25)
        self.dropout2 = nn.Dropout(0.5)
        side = (image_size - 4) // 2
        self.fc1 = nn.Linear(64 * side * side, 128)
        self.fc2 = nn.Linear(128, num_classes)

    def forward(self, x):
        x = F.relu(self.conv1(x))
        x = F.relu(self.conv2(x))
        x = F.max_pool2d(x, 2)
        x = self.dropout1(x)
        x = torch.flatten(x, 1)
        x = F.relu(self.fc1(x))
        x = self.dropout2(x)
        x = self.fc2(x)
        return F.log_softmax(x, dim=1)


def CifarNet() -> BasicNet:
    """``BasicNet`` on CIFAR-10-shaped input (3x32x32)."""
    return BasicNet(in_channels=3, image_size=32, num_classes=10)
