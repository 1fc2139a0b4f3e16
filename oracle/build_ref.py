"""Builds the reference's OWN native component for this path from the sources where they lie under /root/reference
(test / baseline infrastructure only; nothing is copied into the repo):

    python oracle/build_ref.py        ->  oracle/_ref/alt_cuda_corr*.so

VideoFlow/alt_cuda_corr/{correlation.cpp, correlation_kernel.cu} (the reference's setup.py:9-11 passes only -O3 and was
last built for sm_70) compiled with torch.utils.cpp_extension for sm_100: the "existing CUDA" comparator of the
volume-free lookup (BASELINE.md B-altcorr) and a direct pin of oracle.mof.alt_corr_forward.  The output directory is
git-ignored but travels to the GPU box; nothing on the box reads /root/reference.  Skipped quietly when the reference
tree is absent (the GPU box uses the prebuilt file).
"""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
REF = os.environ.get("VFML_REFERENCE", "/root/reference")
SRC = os.path.join(REF, "VideoFlow", "alt_cuda_corr")


def built():
    return os.path.isdir(OUT) and any(f.startswith("alt_cuda_corr") and f.endswith(".so") for f in os.listdir(OUT))


def build(force=False):
    if built() and not force:
        return True
    if not os.path.isfile(os.path.join(SRC, "correlation_kernel.cu")):
        return False
    os.makedirs(OUT, exist_ok=True)
    os.environ.setdefault("TORCH_CUDA_ARCH_LIST", "10.0")
    os.environ.setdefault("MAX_JOBS", "4")
    from torch.utils.cpp_extension import load
    load(name="alt_cuda_corr", sources=[os.path.join(SRC, "correlation.cpp"), os.path.join(SRC, "correlation_kernel.cu")],
         extra_cuda_cflags=["-O3", "-gencode", "arch=compute_100,code=sm_100"], build_directory=OUT, verbose=False,
         is_python_module=False)
    return built()


def load_alt_cuda_corr():
    """Import the prebuilt module (None when it was never built)."""
    if not built():
        return None
    import importlib.util
    import torch  # noqa: F401  (the extension links against libtorch)
    so = [f for f in os.listdir(OUT) if f.startswith("alt_cuda_corr") and f.endswith(".so")][0]
    spec = importlib.util.spec_from_file_location("alt_cuda_corr", os.path.join(OUT, so))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


if __name__ == "__main__":
    ok = build(force="--force" in sys.argv)
    print("oracle/_ref/alt_cuda_corr:", "built" if ok else "reference sources absent, skipped")
