"""VideoFlowInference: back-compat facade, pure delegation (reference processing/flow_inference.py:24-157)."""
from .videoflow_processor import VideoFlowProcessor


class VideoFlowInference:
    def __init__(self, device, fast_mode=False, tile_mode=False, sequence_length=5,
                 dataset='sintel', architecture='mof', variant='standard', precision=None):
        self.device = device
        self.fast_mode = fast_mode
        self.tile_mode = tile_mode
        self.sequence_length = sequence_length
        self.dataset = dataset
        self.architecture = architecture
        self.variant = variant
        self.processor = VideoFlowProcessor(device, fast_mode, tile_mode, sequence_length, dataset, architecture,
                                            variant, precision=precision)
        self.model = None
        self.cfg = None

    def load_model(self):
        self.processor.load_model()
        self.model = self.processor.core.model
        self.cfg = self.processor.core.cfg

    def calculate_tile_grid(self, width, height, tile_size=1280):
        return self.processor.calculate_tile_grid(width, height, tile_size)

    def extract_tile(self, frame, tile_info):
        return self.processor.extract_tile(frame, tile_info)

    def prepare_frame_sequence(self, frames, frame_idx):
        return self.processor.prepare_frame_sequence(frames, frame_idx)

    def compute_optical_flow(self, frames, frame_idx):
        return self.processor.compute_optical_flow(frames, frame_idx)

    def compute_optical_flow_with_progress(self, frames, frame_idx, tile_pbar=None):
        return self.processor.compute_optical_flow_with_progress(frames, frame_idx, tile_pbar)

    def compute_optical_flow_tiled(self, frames, frame_idx, tile_pbar=None, overall_pbar=None):
        return self.processor.compute_optical_flow_tiled(frames, frame_idx, tile_pbar, overall_pbar)

    def is_model_loaded(self):
        return self.processor.is_model_loaded()

    def get_model_info(self):
        return self.processor.get_model_info()

    def get_memory_usage(self):
        return self.processor.get_memory_usage()

    def get_core_engine(self):
        return self.processor.core

    def get_processor(self):
        return self.processor

    def validate_frames(self, frames, frame_idx):
        return self.processor.validate_frames(frames, frame_idx)

    def set_tile_mode(self, enabled):
        self.tile_mode = enabled
        self.processor.set_tile_mode(enabled)

    def set_sequence_length(self, length):
        self.processor.set_sequence_length(length)
        self.sequence_length = length
