"""Shim for `from efficientnet_pytorch import EfficientNet` (orchestration/model_registry.py:16)."""


class EfficientNet:
    @classmethod
    def from_name(cls, name, *args, **kwargs):
        raise RuntimeError(f"efficientnet_pytorch.EfficientNet.from_name('{name}') reached: the b200 backend should "
                           "serve efficientnet_b3 (deepfakedetection_b200.registry.register() not called?)")
