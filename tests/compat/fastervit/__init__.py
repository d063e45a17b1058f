"""Shim for `from fastervit import create_model` (orchestration/model_registry.py:17)."""


def create_model(model_name, *args, **kwargs):
    raise RuntimeError(f"fastervit.create_model('{model_name}') reached: the b200 backend should serve this name "
                       "(deepfakedetection_b200.registry.register() not called?)")
