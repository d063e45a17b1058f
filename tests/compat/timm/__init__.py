"""Shim: lets `orchestration/model_registry.py:15` import.  Constructing a model through it is an error -- the B200
registry entries must have replaced every builder that would call it."""


def create_model(model_name, *args, **kwargs):
    raise RuntimeError(f"timm.create_model('{model_name}') reached: the b200 backend should serve this name "
                       "(deepfakedetection_b200.registry.register() not called?)")
