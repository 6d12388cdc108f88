from ._misc import RegressBoxes, RegressBoxes3D, DenormBoxes3D, UpsampleLike, Anchors, ClipBoxes  # noqa: F401
from .filter_detections import FilterDetections, filter_detections  # noqa: F401
