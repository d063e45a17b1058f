from deepfakedetection_b200.gradcam import show_cam_on_image  # noqa: F401
