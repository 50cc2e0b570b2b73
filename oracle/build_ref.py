"""Compile the reference's OWN CUDA extension (cuda_extension/TriplaneSampler.cpp + TriplaneSampler_kernel.cu),
unmodified and from where the sources lie under /root/reference, into ``oracle/_ref/triplane_sampler_cuda.so``.
TEST / BASELINE INFRASTRUCTURE ONLY: the product never loads it.

It is the second oracle of SURVEY 8a row a9 (the literal drop-in ``enarf_triplane_sample_fwd/bwd``): on the GPU box
``tests/test_gpu_ref_extension.py`` runs the reference's kernels next to ours on the same inputs, and
``scripts/raw_op_bench.py`` times them side by side.  This script is our recipe (torch.utils.cpp_extension + ninja
with an explicit sm_100 target); the reference's setup.py is not run and no reference source is copied -- only the
built ``.so`` lands in ``oracle/_ref/`` (git-ignored, shipped to the GPU box with the snapshot).

    python oracle/build_ref.py
"""
import glob
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
SRC = "/root/reference/cuda_extension"
NAME = "triplane_sampler_cuda"


def so_path():
    hits = glob.glob(os.path.join(OUT, NAME + "*.so"))
    return hits[0] if hits else None


def build(verbose=False):
    """Returns the path of the built module, or None when the reference sources are not present."""
    if so_path():
        return so_path()
    if not os.path.isdir(SRC):
        return None
    from torch.utils import cpp_extension
    work = os.path.join(OUT, "build")
    os.makedirs(work, exist_ok=True)
    os.environ.setdefault("TORCH_CUDA_ARCH_LIST", "10.0")
    os.environ.setdefault("MAX_JOBS", "4")
    cpp_extension.load(name=NAME, sources=[os.path.join(SRC, "TriplaneSampler.cpp"),
                                           os.path.join(SRC, "TriplaneSampler_kernel.cu")],
                       build_directory=work, verbose=verbose, is_python_module=False,
                       extra_cuda_cflags=["-lineinfo"])
    shutil.copy(os.path.join(work, NAME + ".so"), os.path.join(OUT, NAME + ".so"))
    shutil.rmtree(work, ignore_errors=True)
    return so_path()


def load():
    """Import the prebuilt module (needs a CUDA device at call time, not at import)."""
    import importlib.util
    import torch  # noqa: F401  (the module links against libtorch)
    path = so_path()
    if path is None:
        return None
    spec = importlib.util.spec_from_file_location(NAME, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


if __name__ == "__main__":
    print(build(verbose="-v" in sys.argv))
