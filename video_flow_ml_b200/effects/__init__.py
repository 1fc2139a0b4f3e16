from .taa_processor import TAAProcessor

__all__ = ["TAAProcessor"]
