"""Shim for `from pytorch_grad_cam import GradCAM` (web_ui.py:22): the backend's pytorch-grad-cam compatible front."""
from deepfakedetection_b200.gradcam import ClassifierOutputTarget, GradCAM, show_cam_on_image  # noqa: F401
