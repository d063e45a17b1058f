from deepfakedetection_b200.gradcam import ClassifierOutputTarget  # noqa: F401
