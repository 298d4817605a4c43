from .camera_head import CameraHead

__all__ = ["CameraHead"]
