"""Build recipe for the reference's own CUDA extension (test infrastructure, not product).

Compiles the UNMODIFIED sources of the reference's rasterizer where they lie
(`/root/reference/submodules/GStex_cuda/gstex_cuda/cuda/csrc/*.cu|*.cpp`) for sm_100 with the
same flags the reference's JIT loader uses (`gstex_cuda/cuda/_backend.py:38-40`: `-O3` only), and
puts the result in `oracle/_ref/gstex_ref_C.so` (git-ignored, travels to the GPU box).

The built module is the *GPU oracle* ("tier T2" in SURVEY.md section 8c): the `-m gpu` tests compare
our kernels with it bit-for-bit (binning, sort, indices) and within tolerance (images, gradients),
and `bench.py` times it beside our kernels as "the reference kernels on the same B200".
No reference source is copied into this repository.

Run:  python oracle/build_ref.py        (about 7 minutes on 8 cores; no GPU needed)
"""
import glob
import os
import shutil
import sys

REF_CSRC = "/root/reference/submodules/GStex_cuda/gstex_cuda/cuda/csrc"
HERE = os.path.dirname(os.path.abspath(__file__))
OUT_DIR = os.path.join(HERE, "_ref")
NAME = "gstex_ref_C"


def build(verbose: bool = True) -> str:
    out = os.path.join(OUT_DIR, NAME + ".so")
    if os.path.exists(out):
        return out
    if not os.path.isdir(REF_CSRC):
        raise FileNotFoundError(f"reference sources not present at {REF_CSRC}")
    os.makedirs(OUT_DIR, exist_ok=True)
    build_dir = os.path.join(OUT_DIR, "build")
    os.makedirs(build_dir, exist_ok=True)
    try:
        os.remove(os.path.join(build_dir, "lock"))
    except OSError:
        pass
    # +PTX so the contraction pattern nvcc chose for get_aabb_2d_kernel can be read back with
    # `cuobjdump -ptx`; the SASS that runs is the sm_100 one either way.
    os.environ["TORCH_CUDA_ARCH_LIST"] = "10.0+PTX"
    os.environ.setdefault("MAX_JOBS", str(os.cpu_count() or 8))
    from torch.utils.cpp_extension import load

    sources = sorted(glob.glob(os.path.join(REF_CSRC, "*.cu"))) + sorted(
        glob.glob(os.path.join(REF_CSRC, "*.cpp"))
    )
    load(
        name=NAME,
        sources=sources,
        extra_cflags=["-O3"],
        extra_cuda_cflags=["-O3"],
        extra_include_paths=[os.path.join(REF_CSRC, "third_party/glm")],
        build_directory=build_dir,
        verbose=verbose,
    )
    shutil.copy2(os.path.join(build_dir, NAME + ".so"), out)
    return out


if __name__ == "__main__":
    print(build(verbose="-q" not in sys.argv))
