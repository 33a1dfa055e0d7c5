"""Model registry (reference ``src/config.py:7``: ``create_network`` keyed by ``M``)."""
from .basicnet import BasicNet, CifarNet
from .googlenet import GoogLeNet, GoogLeNetBackbone
from .resnet import ResNet50, ResNetBackbone

create_network = {
    "mnist": BasicNet,
    "cifar": CifarNet,
    "imagenet": GoogLeNet,
    "googlenet": GoogLeNet,
    "resnet50": ResNet50,
}

__all__ = ["BasicNet", "CifarNet", "GoogLeNet", "GoogLeNetBackbone", "ResNet50", "ResNetBackbone", "create_network"]
