"""Compiles the reference's OWN CUDA extensions (unmodified sources, read in place from /root/reference/op) for sm_100.

TEST INFRASTRUCTURE ONLY.  Output goes to oracle/_ref/<name>/<name>.so (git-ignored, but it travels to the GPU box
with the repo snapshot).  On the GPU box these are the second oracle ("does our kernel agree with the reference's
kernel on the same inputs") and the "kernel to beat" in bench scripts.  Nothing is copied from the reference: nvcc
reads the sources where they lie.

    python oracle/build_ref_ext.py [upfirdn2d] [fused] [ns_step_forward] [gridsample_grad2]
"""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF_OP = os.path.join(os.environ.get("B200_REFERENCE_DIR", "/root/reference"), "op")
OUT = os.path.join(HERE, "_ref")

EXTS = {
    "upfirdn2d": ["upfirdn2d.cpp", "upfirdn2d_kernel.cu"],            # op/upfirdn2d.py:10-16
    "fused": ["fused_bias_act.cpp", "fused_bias_act_kernel.cu"],      # op/fused_act.py:11-17
    "ns_step_forward": ["ns_step.cpp", "ns_step_kernel.cu"],          # op/ns_step.py:11-17
    "gridsample_grad2": ["grid_sample.cpp", "grid_sample_kernel.cu"]  # op/grid_sample.py:8-13
}


def build(names):
    os.environ.setdefault("TORCH_CUDA_ARCH_LIST", "10.0")
    os.environ.setdefault("MAX_JOBS", "4")
    from torch.utils.cpp_extension import load
    for n in names:
        d = os.path.join(OUT, n)
        os.makedirs(d, exist_ok=True)
        if os.path.exists(os.path.join(d, n + ".so")):
            print("have", n)
            continue
        print("building", n, flush=True)
        load(n, sources=[os.path.join(REF_OP, s) for s in EXTS[n]], build_directory=d, verbose=False)
        print("built", n, flush=True)


def load_ext(name):
    """Import a previously built extension (works on the GPU box, where /root/reference is absent)."""
    import importlib.util
    import torch  # noqa: F401  (the extension links against libtorch)
    path = os.path.join(OUT, name, name + ".so")
    if not os.path.exists(path):
        return None
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


if __name__ == "__main__":
    build(sys.argv[1:] or ["upfirdn2d", "fused", "ns_step_forward"])
