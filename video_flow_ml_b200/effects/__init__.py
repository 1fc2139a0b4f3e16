from .taa_processor import TAAComparisonProcessor, TAAProcessor, apply_taa_effect

__all__ = ["TAAProcessor", "TAAComparisonProcessor", "apply_taa_effect"]
