"""deepfakedetection_b200 -- B200-native (sm_100a) engine for the DeepfakeDetection eval hot path.

Public surface:
  * ``registry.register()``      -- plug the backend into the reference's model registry
  * ``registry.get_model_spec``  -- standalone lookup with the same semantics
  * ``module.B200EfficientNetB0``-- the nn.Module the builder returns
  * ``gradcam.GradCAM``          -- pytorch-grad-cam compatible front of the fused CAM kernel
  * ``evaluate``                 -- batch-sharded multi-GPU evaluation harness
  * ``_lib``                     -- ctypes binding of libb200df.so (C ABI: include/b200df.h)
"""

__all__ = ["registry", "module", "gradcam", "evaluate", "_lib"]
__version__ = "0.1.0"
