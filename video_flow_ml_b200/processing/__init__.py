"""Drop-in mirror of the reference's ``processing`` package (processing/__init__.py:23-40) for the
VideoFlow path: VideoFlowInference -> VideoFlowProcessor -> VideoFlowCore."""
from .flow_inference import VideoFlowInference
from .videoflow_core import VideoFlowCore
from .videoflow_processor import VideoFlowProcessor

__all__ = ["VideoFlowInference", "VideoFlowCore", "VideoFlowProcessor"]
