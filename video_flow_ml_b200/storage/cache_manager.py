"""Flow cache formats and the LOD pyramid: mirrors storage/cache_manager.py of the reference
(FlowFileHandler :14-70, LODGenerator :73-161, the per-frame load / save / LOD methods of FlowCacheManager :221-337).

The file formats are byte-for-byte the reference's (Middlebury ``.flo``: 'PIEH', width, height as little-endian u32,
fp32 data; ``.npz``: ``np.savez_compressed`` with the same keys).  The LOD pyramid -- a pure-Python per-pixel loop in the
reference, its slowest host step after inference -- is one libvfml kernel launch (``vfml_lod_pyramid``) and is
bit-identical with that loop.  Flows may be numpy arrays or CUDA tensors (a device flow never visits the host for the
reduction; it is downloaded once per level only when a file is written).  Cache-directory naming
(``generate_cache_path`` -> storage/filename_generator.py) is application logic and not part of this build."""
import os
import struct

import numpy as np
import torch

from .. import ops


def _to_numpy(flow):
    return flow.detach().cpu().numpy() if torch.is_tensor(flow) else flow


class FlowFileHandler:
    @staticmethod
    def save_flow_flo(flow, filename):
        flow = _to_numpy(flow)
        height, width = flow.shape[:2]
        with open(filename, 'wb') as f:
            f.write(b'PIEH')
            f.write(struct.pack('<I', width))
            f.write(struct.pack('<I', height))
            f.write(flow.astype(np.float32).tobytes())

    @staticmethod
    def save_flow_npz(flow, filename, frame_idx=None, metadata=None):
        save_data = {'flow': _to_numpy(flow).astype(np.float32)}
        if frame_idx is not None:
            save_data['frame_idx'] = frame_idx
        if metadata is not None:
            save_data.update(metadata)
        np.savez_compressed(filename, **save_data)

    @staticmethod
    def load_flow_flo(filename):
        with open(filename, 'rb') as f:
            magic = f.read(4)
            if magic != b'PIEH':
                raise ValueError(f"Invalid .flo file magic number: {magic}")
            width = struct.unpack('<I', f.read(4))[0]
            height = struct.unpack('<I', f.read(4))[0]
            data = f.read(width * height * 2 * 4)
        return np.frombuffer(data, dtype=np.float32).reshape(height, width, 2)

    @staticmethod
    def load_flow_npz(filename):
        return dict(np.load(filename))


class LODGenerator:
    @staticmethod
    def generate_lods(flow, num_lods=5, device="cuda"):
        """[original, lod1, ...]: each level halves the size (rounding up) by averaging the existing cells of every
        2x2 block and halving the vectors.  numpy in -> numpy out (the reference's types); CUDA tensor in -> CUDA
        tensors out."""
        if torch.is_tensor(flow) and flow.is_cuda:
            return ops.lod_pyramid(flow.contiguous().float(), num_lods)
        arr = _to_numpy(flow)
        t = torch.from_numpy(np.ascontiguousarray(arr, dtype=np.float32)).to(device)
        lods = ops.lod_pyramid(t, num_lods)
        return [arr] + [l.cpu().numpy() for l in lods[1:]]


class FlowCacheManager:
    def __init__(self):
        self.file_handler = FlowFileHandler()
        self.lod_generator = LODGenerator()

    @staticmethod
    def _frame_file(cache_dir, frame_idx, ext):
        return os.path.join(cache_dir, f"flow_frame_{frame_idx:06d}.{ext}")

    def check_cache_exists(self, cache_dir, max_frames):
        if not os.path.exists(cache_dir):
            return False, None, list(range(max_frames))
        files = os.listdir(cache_dir)
        if any(f.endswith('.npz') for f in files):
            format_type = 'npz'
        elif any(f.endswith('.flo') for f in files):
            format_type = 'flo'
        else:
            return False, None, list(range(max_frames))
        missing = [i for i in range(max_frames) if not os.path.exists(self._frame_file(cache_dir, i, format_type))]
        return (not missing), (format_type if not missing else None), missing

    def load_cached_flow(self, cache_dir, frame_idx, format_type='auto'):
        npz_file, flo_file = self._frame_file(cache_dir, frame_idx, "npz"), self._frame_file(cache_dir, frame_idx, "flo")
        if format_type == 'auto':
            if os.path.exists(npz_file):
                return self.file_handler.load_flow_npz(npz_file)['flow']
            if os.path.exists(flo_file):
                return self.file_handler.load_flow_flo(flo_file)
            raise FileNotFoundError(f"No cached flow found for frame {frame_idx}")
        if format_type == 'npz':
            return self.file_handler.load_flow_npz(npz_file)['flow']
        if format_type == 'flo':
            return self.file_handler.load_flow_flo(flo_file)
        raise ValueError(f"Invalid format_type: {format_type}")

    def save_flow_to_cache(self, flow, cache_dir, frame_idx, save_format='npz'):
        os.makedirs(cache_dir, exist_ok=True)
        flow = _to_numpy(flow)
        if save_format in ['flo', 'both']:
            self.file_handler.save_flow_flo(flow, self._frame_file(cache_dir, frame_idx, "flo"))
        if save_format in ['npz', 'both']:
            metadata = {'frame_idx': frame_idx, 'shape': flow.shape, 'dtype': str(flow.dtype)}
            self.file_handler.save_flow_npz(flow, self._frame_file(cache_dir, frame_idx, "npz"), frame_idx, metadata)

    def save_flow_lods(self, lods, cache_dir, frame_idx):
        os.makedirs(cache_dir, exist_ok=True)
        for lod_level, lod_data in enumerate(lods):
            lod_data = _to_numpy(lod_data)
            filename = os.path.join(cache_dir, f"flow_frame_{frame_idx:06d}_lod{lod_level}.npz")
            metadata = {'frame_idx': frame_idx, 'lod_level': lod_level, 'shape': lod_data.shape, 'dtype': str(lod_data.dtype)}
            self.file_handler.save_flow_npz(lod_data, filename, frame_idx, metadata)

    def load_flow_lod(self, cache_dir, frame_idx, lod_level=0):
        filename = os.path.join(cache_dir, f"flow_frame_{frame_idx:06d}_lod{lod_level}.npz")
        if not os.path.exists(filename):
            raise FileNotFoundError(f"LOD {lod_level} not found for frame {frame_idx}")
        return self.file_handler.load_flow_npz(filename)['flow']

    def check_flow_lods_exist(self, cache_dir, max_frames, num_lods=5):
        if not os.path.exists(cache_dir):
            return False
        return all(os.path.exists(os.path.join(cache_dir, f"flow_frame_{i:06d}_lod{l}.npz"))
                   for i in range(max_frames) for l in range(num_lods))

    def generate_lods_for_cache(self, cache_dir, max_frames, num_lods=5):
        for frame_idx in range(max_frames):
            lods = self.lod_generator.generate_lods(self.load_cached_flow(cache_dir, frame_idx), num_lods)
            self.save_flow_lods(lods, cache_dir, frame_idx)
