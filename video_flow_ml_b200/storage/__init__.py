from .cache_manager import FlowCacheManager, FlowFileHandler, LODGenerator

__all__ = ["FlowFileHandler", "LODGenerator", "FlowCacheManager"]
