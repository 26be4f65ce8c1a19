#!/usr/bin/env python
"""Install the UNMODIFIED reference package where the GPU box can see it: /root/reference/vllm_i64 -> oracle/_ref/vllm_i64.

oracle/_ref/ is git-ignored (the history stays source-only) but NOT gpurun-ignored, so it travels to the GPU box with the
snapshot exactly like the built libi64b.so.  It is test / bench infrastructure only (the product never imports it):

  * bench.py --impl reference       : the reference's own TokenRoutedMLP -> _fuse_cpu_projection_layers ->
                                      _quantize_cpu_dynamic (core/loader.py:57-94) timed on the host cores
                                      (cpu_baseline.kind = "reference"), and CPUEngine for config 1;
  * bench.py --impl reference-gpu   : the reference's TokenRoutedMLP.forward on cuda bf16 (the "before" number);
  * tests/test_dropin_reference.py  : the reference's ComplexityDeepModel built over this repo's MLPs.

/root/reference is read-only and exists only in the build container; on the GPU box this script is a no-op and the
callers fall back (port / skip) when oracle/_ref is missing.  Run by __graft_entry__.build().
"""
import os
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.environ.get("I64B_REFERENCE_ROOT", "/root/reference")
DST = os.path.join(ROOT, "oracle", "_ref")


def install(verbose: bool = True) -> bool:
    src_pkg = os.path.join(SRC, "vllm_i64")
    if not os.path.isdir(src_pkg):
        if verbose:
            print(f"install_ref: {src_pkg} not present (GPU box?): keeping whatever is in {DST}")
        return os.path.isdir(os.path.join(DST, "vllm_i64"))
    dst_pkg = os.path.join(DST, "vllm_i64")
    if os.path.isdir(dst_pkg):
        shutil.rmtree(dst_pkg)
    shutil.copytree(src_pkg, dst_pkg, ignore=shutil.ignore_patterns("__pycache__", "*.pyc"))
    with open(os.path.join(DST, "INSTALLED_FROM"), "w") as f:
        f.write(f"{SRC} (unmodified copy made by tools/install_ref.py; not tracked by git)\n")
    if verbose:
        n = sum(len(fs) for _, _, fs in os.walk(dst_pkg))
        print(f"install_ref: copied {n} files to {dst_pkg}")
    return True


if __name__ == "__main__":
    sys.exit(0 if install() else 1)
