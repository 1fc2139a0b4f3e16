"""Fixtures for the remaining encoders / decoders, the LOD pyramid and the quality map, every array
an output of the UNMODIFIED reference:   python tests/golden/make_golden_post2.py  -> post2.npz

  encoding/flow_encoders.py   MotionVectorsRGB8FlowEncoder.encode / .decode ('rgb+'), HSVFlowEncoder.encode,
                              TorchvisionFlowEncoder.encode, encode_motion_vectors / decode_motion_vectors defaults
  storage/cache_manager.py    FlowFileHandler.save_flow_flo (bytes), LODGenerator.generate_lods
  correction_worker.py        generate_quality_frame_gpu (run on 'cpu')
"""
import importlib.util
import os
import sys
import tempfile
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import cases      # noqa: E402

REF = os.environ.get("VFML_REFERENCE", "/root/reference")
sys.path.insert(0, REF)


def load_functions(path, names):
    """correction_worker.py imports tkinter-free modules only, but executes nothing at import besides defs; load it
    with stubs for the packages that are absent here."""
    for m in ("cv2",):
        try:
            __import__(m)
        except ImportError:
            sys.modules[m] = types.ModuleType(m)
    spec = importlib.util.spec_from_file_location("ref_" + os.path.basename(path)[:-3], path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return [getattr(mod, n) for n in names]


def main():
    import encoding.flow_encoders as fe
    import storage.cache_manager as cm
    out = {}
    ff = cases.flow_field_case()
    h, w = ff.shape[:2]
    with np.errstate(all="ignore"):
        out["rgb8"] = fe.MotionVectorsRGB8FlowEncoder().encode(ff.copy(), w, h)
        out["rgb8_c8"] = fe.MotionVectorsRGB8FlowEncoder(clamp_range=8.0).encode(ff.copy(), w, h)
        out["rgb8_decoded"] = fe.MotionVectorsRGB8FlowEncoder().decode(out["rgb8"])
        out["mv_default"] = fe.encode_motion_vectors(ff.copy())                      # format_variant defaults to 'rgb8', clamp 64
        out["mv_default_decoded"] = fe.decode_motion_vectors(out["mv_default"])
        out["hsv"] = fe.HSVFlowEncoder().encode(ff.copy(), w, h)
        out["hsv_zero"] = fe.HSVFlowEncoder().encode(np.zeros((5, 7, 2), np.float32), 7, 5)
        fin = cases.flow_field_case(seed=31, h=45, w=61)
        fin = np.nan_to_num(fin, nan=0.5, posinf=3.0, neginf=-3.0)                   # torchvision: finite input
        out["tv"] = fe.TorchvisionFlowEncoder().encode(fin.copy(), 61, 45)
        out["hsv_fin"] = fe.encode_flow(fin.copy(), 61, 45, "hsv")
    with tempfile.TemporaryDirectory() as tmp:
        p = os.path.join(tmp, "a.flo")
        cm.FlowFileHandler.save_flow_flo(fin, p)
        out["flo_bytes"] = np.frombuffer(open(p, "rb").read(), np.uint8)
        assert np.array_equal(cm.FlowFileHandler.load_flow_flo(p), fin)
    for name, f in (("lod_odd", fin), ("lod_even", np.nan_to_num(cases.flow_field_case(seed=37, h=64, w=96), nan=1.0, posinf=2.0, neginf=-2.0))):
        for i, l in enumerate(cm.LODGenerator.generate_lods(f, 5)):
            if i:
                out[f"{name}_{i}"] = l
    (qf,) = load_functions(os.path.join(REF, "correction_worker.py"), ["generate_quality_frame_gpu"])
    frames, flows = cases.taa_case(seed=41, n=3, h=33, w=47)
    fl = flows[0] * 2
    fl[3, 3] = [100.0, 0.0]
    fl[4, 4] = [0.0, -100.0]
    out["quality"] = qf(frames[0], frames[1], fl, "cpu", 0.7)
    out["quality_same"] = qf(frames[0], frames[0], np.zeros_like(fl), "cpu", 0.7)
    out["quality_resized"] = qf(frames[0], frames[1], np.ascontiguousarray(fl[::2, ::2][:16, :23]), "cpu", 0.7)
    np.savez_compressed(os.path.join(HERE, "post2.npz"), **out)
    print("wrote post2.npz", {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
