"""Offline stand-in for the torch.hub checkout of pytorch/vision v0.6.0.

The unmodified reference builds its ImageNet model with
``torch.hub.load('pytorch/vision:v0.6.0', 'googlenet', pretrained=False, init_weights=False)``
(reference src/network.py:36), which downloads a GitHub archive.  There is no network on the
build or GPU boxes, so bench.py points ``TORCH_HOME`` at this directory; torch.hub then finds
"pytorch_vision_v0.6.0" in its cache and calls the entry point below, which returns the same
architecture from the installed torchvision (aux heads on, no input transform, default init).
Nothing in the reference itself is changed.
"""
dependencies = ["torch", "torchvision"]


def googlenet(pretrained=False, progress=True, **kwargs):
    from torchvision.models import googlenet as _googlenet

    if pretrained:
        raise RuntimeError("offline hub stand-in: pretrained weights are not available")
    kwargs.setdefault("aux_logits", True)
    kwargs.setdefault("transform_input", False)
    return _googlenet(weights=None, **kwargs)
