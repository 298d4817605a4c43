"""API mirrors of sailrecon/utils/{pose_enc,rotation,geometry}.py for the functions on the path right after the camera
head."""
from .geometry import unproject_depth_map_to_point_map  # noqa: F401
from .pose_enc import pose_encoding_to_extri_intri  # noqa: F401
from .rotation import quat_to_mat  # noqa: F401
