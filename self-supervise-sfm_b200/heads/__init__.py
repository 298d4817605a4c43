from .camera_head import CameraHead
from .dpt_head import DPTHead

__all__ = ["CameraHead", "DPTHead"]
