"""Make the UNMODIFIED reference travel to the GPU box.  TEST / BASELINE INFRASTRUCTURE ONLY: the product never
loads anything from ``oracle/_ref``.

1. Compile the reference's OWN CUDA extension (cuda_extension/TriplaneSampler.cpp + TriplaneSampler_kernel.cu),
   unmodified and from where the sources lie under /root/reference, into ``oracle/_ref/triplane_sampler_cuda.so``.
   It is the second oracle of SURVEY 8a row a9 (the literal drop-in ``enarf_triplane_sample_fwd/bwd``): on the GPU box
   ``tests/test_gpu_ref_extension.py`` runs the reference's kernels next to ours on the same inputs, and
   ``scripts/raw_op_bench.py`` times them side by side.  This script is our recipe (torch.utils.cpp_extension + ninja
   with an explicit sm_100 target); the reference's setup.py is not run.
2. Pack the reference's hot-path Python (the files ``models.generator`` imports, the StyleGAN2 discriminator /
   tri-plane generator they instantiate and the GAN losses) into ``oracle/_ref/reference_py.zip`` -- a build output of
   this recipe, git-ignored like the ``.so`` (the repository history holds no reference source), shipped to the GPU
   box with the snapshot.  ``tests/golden/ref_harness.py`` imports the reference from ``/root/reference`` when that
   exists and from the archive otherwise, so on the B200 the reference's eager ``grid_sample`` path runs on CUDA as
   the second oracle, as the ``--impl reference`` arm and as the GPU baseline of BASELINE.md.

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


# ---- the reference's Python hot path ----------------------------------------------------------------------
REF = "/root/reference"
PY_ZIP = os.path.join(OUT, "reference_py.zip")
# directories are taken whole (``*.py`` only), single files as they are; everything else of the reference
# (datasets, training scripts, evaluation, mmpose, TF StyleGAN2) stays where it is
PY_TREES = ["models", "libraries/NeRF", "libraries/NARF", "libraries/triplane", "libraries/custom_stylegan2",
            "libraries/gan", "libraries/stylegan2_ada_pytorch/dnnlib", "libraries/stylegan2_ada_pytorch/torch_utils",
            "libraries/stylegan2_ada_pytorch/training"]
PY_FILES = ["libraries/pytorch3d_utils.py", "libraries/stylegan2_pytorch/model.py", "cuda_extension/triplane_sampler.py"]


def py_root():
    """Import root of the reference: the live tree in the authoring container, else the staged archive, else None."""
    if os.path.isdir(os.path.join(REF, "libraries")):
        return REF
    return PY_ZIP if os.path.exists(PY_ZIP) else None


def stage_python(force=False):
    """Pack the hot-path ``*.py`` files of /root/reference, byte-identical, into one archive oracle/_ref/reference_py.zip
    (Python imports straight from it: no extracted copy of the reference ever lies in the repository tree)."""
    import zipfile
    if os.path.exists(PY_ZIP) and not force:
        return PY_ZIP
    if not os.path.isdir(REF):
        return None
    os.makedirs(OUT, exist_ok=True)
    rels = list(PY_FILES)
    for tree in PY_TREES:
        for d, _, files in os.walk(os.path.join(REF, tree)):
            rels += [os.path.relpath(os.path.join(d, f), REF) for f in files if f.endswith(".py")]
    rels = sorted(set(rels))
    # explicit directory entries: the reference's packages have no __init__.py, and zipimport only treats a
    # directory as a namespace-package portion when the archive lists it
    dirs = sorted({rel[:i] + "/" for rel in rels for i, ch in enumerate(rel) if ch == "/"})
    with zipfile.ZipFile(PY_ZIP, "w", zipfile.ZIP_DEFLATED) as z:
        for d in dirs:
            z.writestr(zipfile.ZipInfo(d), "")
        for rel in rels:
            z.write(os.path.join(REF, rel), rel)
    return PY_ZIP


if __name__ == "__main__":
    print(stage_python(force="--force" in sys.argv))
    if "--py-only" not in sys.argv:
        print(build(verbose="-v" in sys.argv))
