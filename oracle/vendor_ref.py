#!/usr/bin/env python
"""Stage the reference's own files for the GPU box (TEST / BASELINE INFRASTRUCTURE).

/root/reference exists only in the build container.  The reference is three BSD-3 Python files, so its "build" is
a copy: this recipe places them, unmodified, in oracle/_ref/ (git-ignored, so no reference source enters the
history; not gpurun-ignored, so they travel with the snapshot like the built .so).  With them on the box

* `bench.py --impl reference` / `cpu_baseline` time the reference's own VAE / loss_function / utils on the host
  cores (`kind: "reference"`) instead of the torch port;
* `bench.py` can time the unchanged reference step on torch-CUDA eager + NCCL (`torch_cuda_eager`);
* `tests/test_gpu_reference_script.py` runs the unchanged vae-hpo.py through the drop-in.

    python oracle/vendor_ref.py            # called by __graft_entry__.build() when /root/reference is present
"""
import filecmp
import os
import shutil
import sys

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
DST = os.path.join(HERE, "_ref")
FILES = ("utils.py", "vae-hpo.py", "example-subgroup.py", "LICENSE", "README.md")


def vendor(verbose=True):
    if not os.path.isdir(REF):
        if verbose:
            print(f"[vendor_ref] {REF} not present: keeping whatever {DST} holds")
        return False
    os.makedirs(DST, exist_ok=True)
    for f in FILES:
        src, dst = os.path.join(REF, f), os.path.join(DST, f)
        if os.path.exists(src) and not (os.path.exists(dst) and filecmp.cmp(src, dst, shallow=False)):
            shutil.copyfile(src, dst)
    if verbose:
        print(f"[vendor_ref] {', '.join(FILES)} -> {DST}")
    return True


if __name__ == "__main__":
    sys.exit(0 if vendor() or os.path.isdir(DST) else 1)
