"""Drop-in mirror of the reference's ``processing`` package (processing/__init__.py:23-40):
VideoFlowInference -> VideoFlowProcessor -> VideoFlowCore and MemFlowInference -> MemFlowProcessor -> MemFlowCore."""
from .flow_inference import VideoFlowInference
from .videoflow_core import VideoFlowCore
from .videoflow_processor import VideoFlowProcessor
from .memflow_core import InferenceCore, MemFlowCore
from .memflow_processor import MemFlowInference, MemFlowProcessor

__all__ = ["VideoFlowInference", "VideoFlowCore", "VideoFlowProcessor",
           "MemFlowInference", "MemFlowCore", "MemFlowProcessor", "InferenceCore"]
