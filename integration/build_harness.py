#!/usr/bin/env python3
"""Compiles the Siril-side glue (integration/siril/src/gpu/gpu_stack.c) against the REFERENCE's own
src/stacking/stacking.h (read where it lies under /root/reference, never copied), the harness's
stand-ins for the rest of Siril (integration/harness) and libsgpu.so, into
integration/_build/libgpu_stack_harness.so.  Needs /root/reference and oracle/_ref/gen (build_ref.py)."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = "/root/reference/src"


def build():
    if not os.path.isdir(REF):
        raise SystemExit("the reference sources are not here: the prebuilt harness library is used as it is")
    gen = os.path.join(ROOT, "oracle", "_ref")
    if not os.path.isdir(os.path.join(gen, "gen")):
        subprocess.check_call([sys.executable, os.path.join(ROOT, "oracle", "build_ref.py")], stdout=subprocess.DEVNULL)
    out = os.path.join(HERE, "_build")
    os.makedirs(out, exist_ok=True)
    so = os.path.join(out, "libgpu_stack_harness.so")
    # the glue says "gpu/sgpu.h" and "gpu/gpu_stack.h", as it would inside a Siril tree (src/gpu)
    inc = os.path.join(out, "include", "gpu")
    os.makedirs(inc, exist_ok=True)
    for src, name in ((os.path.join(ROOT, "include", "sgpu.h"), "sgpu.h"),
                      (os.path.join(HERE, "siril", "src", "gpu", "gpu_stack.h"), "gpu_stack.h")):
        dst = os.path.join(inc, name)
        if os.path.islink(dst) or os.path.exists(dst):
            os.remove(dst)
        os.symlink(src, dst)
    cmd = ["gcc", "-O2", "-g", "-std=gnu11", "-Wall", "-Wno-unused-variable", "-fPIC", "-shared",
           "-I", os.path.join(HERE, "harness", "include"), "-I", os.path.join(out, "include"), "-I", gen, "-I", REF,
           os.path.join(HERE, "siril", "src", "gpu", "gpu_stack.c"), os.path.join(HERE, "harness", "harness.c"),
           os.path.join(HERE, "harness", "tdist.c"),
           "-L", os.path.join(ROOT, "siril_b200"), "-lsgpu", "-Wl,-rpath,$ORIGIN/../../siril_b200", "-lm", "-o", so]
    subprocess.check_call(cmd)
    return so


if __name__ == "__main__":
    print(build())
