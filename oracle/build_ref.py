#!/usr/bin/env python3
"""TEST INFRASTRUCTURE ONLY — builds oracle/_ref/libsiril_ref.so from the *reference's own
sources where they lie* under /root/reference (never copied into the repo).

Two kinds of input:
  * whole, unmodified translation units compiled with shim headers
    (oracle/shim precedes the reference include path):
        src/stacking/rejection_float.c, src/stacking/siril_fit_linear.c,
        src/algos/sorting.c, src/rt/rt_algo.cc
  * line ranges of TUs that are too entangled with glib/cfitsio/opencv to compile
    whole; they are extracted at build time into oracle/_ref/gen/*.inc (git-ignored)
    and #included by oracle/ref_driver.c.  Each range is pinned by SHA-256 so that a
    changed reference fails loudly instead of silently extracting the wrong lines.

Outputs go only to oracle/_ref/ (git-ignored, NOT gpurun-ignored: the .so travels to the
GPU box, /root/reference does not).

Usage: python oracle/build_ref.py [--update-hashes]
"""
import hashlib
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("SIRIL_REFERENCE", "/root/reference")
OUT = os.path.join(HERE, "_ref")
GEN = os.path.join(OUT, "gen")

# name -> (reference file, first line, last line, sha256 of the extracted text)
RANGES = {
    # --- shim content (types / macros / inlines whose exact text matters) ---
    "siril_h_macros":    ("src/core/siril.h", 60, 81, None),
    "siril_h_consts":    ("src/core/siril.h", 88, 99, None),
    "siril_h_threading": ("src/core/siril.h", 359, 363, None),
    "siril_h_imstats":   ("src/core/siril.h", 406, 424, None),
    "settings_enums":    ("src/core/settings.h", 24, 52, None),
    "proto_round_to_int":   ("src/core/proto.h", 208, 213, None),
    "proto_round_to_WORD":  ("src/core/proto.h", 232, 237, None),
    "proto_ceil_multiple":  ("src/core/proto.h", 295, 299, None),
    "proto_truncate_WORD":  ("src/core/proto.h", 324, 328, None),
    "proto_roundf_to_WORD": ("src/core/proto.h", 341, 346, None),
    "proto_set_float_in_interval": ("src/core/proto.h", 384, 388, None),
    "proto_double_ushort_to_float": ("src/core/proto.h", 418, 420, None),
    # --- hot-path code ranges ---
    "mm_planner":        ("src/stacking/median_and_mean.c", 261, 356, None),
    "mm_norm01":         ("src/stacking/median_and_mean.c", 557, 582, None),
    "mm_rejection_u16":  ("src/stacking/median_and_mean.c", 590, 1101, None),
    "mm_noise_weights":  ("src/stacking/median_and_mean.c", 1111, 1135, None),
    "mm_maxframe_accum": ("src/stacking/median_and_mean.c", 161, 169, None),
    "mm_maxframe_final": ("src/stacking/median_and_mean.c", 182, 189, None),
    "mm_other_weights":  ("src/stacking/median_and_mean.c", 1137, 1230, None),
    "mm_linfit_precalc": ("src/stacking/median_and_mean.c", 1491, 1499, None),
    "mm_mask_ramp":      ("src/stacking/median_and_mean.c", 513, 521, None),
    "blend_ramp":        ("src/stacking/blending.c", 30, 50, None),
    "stats_u16_sd_mad":  ("src/algos/statistics.c", 100, 154, None),
    "statsf_mad_bwmv":   ("src/algos/statistics_float.c", 79, 127, None),
    "statsf_ikss":       ("src/algos/statistics_float.c", 199, 229, None),
    "quant_defs":        ("src/algos/quantize.c", 38, 39, None),
    "quant_meansigma":   ("src/algos/quantize.c", 207, 317, None),
    "quant_diffmeansigma": ("src/algos/quantize.c", 327, 422, None),
    "quant_noise1":      ("src/algos/quantize.c", 1202, 1501, None),
    "norm_factors":      ("src/stacking/normalization.c", 144, 181, None),
    # --- overlap normalization ---
    "norm_solve_overlap":   ("src/stacking/normalization.c", 296, 355, None),
    "norm_compute_overlap": ("src/stacking/normalization.c", 420, 456, None),
    "norm_overlap_layers":  ("src/stacking/normalization.c", 549, 593, None),
    "norm_overlap_solves":  ("src/stacking/normalization.c", 875, 906, None),
    # --- statistics records of the .seq file ---
    "seq_stats_write":   ("src/io/seqfile.c", 838, 853, None),
    "seq_stats_read":    ("src/io/seqfile.c", 584, 600, None),
    # --- SER frames ---
    "ser_enums":         ("src/io/ser.h", 70, 85, None),
    "ser_endian_depth":  ("src/io/ser.c", 521, 539, None),
    "reg_translation_from_H": ("src/registration/registration.c", 301, 304, None),
    # --- sum / min / max stackers (ref_simple.c) ---
    "sum_pixel_loop":    ("src/stacking/sum.c", 186, 231, None),
    "sum_find_max":      ("src/stacking/sum.c", 257, 279, None),
    "sum_convert":       ("src/stacking/sum.c", 308, 344, None),
    "minmax_pixel_loop": ("src/stacking/stackminmax.c", 160, 182, None),
}

HASHES_FILE = os.path.join(HERE, "ref_ranges.sha256")

WHOLE_TUS = [
    ("src/stacking/rejection_float.c", "gcc"),
    ("src/stacking/siril_fit_linear.c", "gcc"),
    ("src/algos/sorting.c", "gcc"),
    ("src/rt/rt_algo.cc", "g++"),
]

# reference default build class: debugoptimized (-O2) + -fno-math-errno (meson.build:36,159).
# Two variants are built:
#  libsiril_ref.so      "pinned order": the reference TUs are compiled WITHOUT -fopenmp.  The
#      reference marks four per-stack reductions `#pragma omp simd reduction(+)`
#      (statistics.h:94-102, rejection_float.c:270, median_and_mean.c:1026,1087), which makes
#      their summation ORDER compiler- and ISA-defined (last double ulp of a sum, last float ulp
#      of the linear-fit sigma).  Parity is pinned on the one order the C abstract machine
#      defines — the sequential one, which is also what the reference's own scalar branches
#      (N < 16 / N < 24) and any non-OpenMP build produce.  Row-block threading comes from
#      ref_omp.c, the only TU compiled with -fopenmp.
#  libsiril_ref_fast.so  everything with -fopenmp as the reference's meson build does; used only
#      as the TIMED CPU baseline (bench.py), never for parity.
BASE = ["-O2", "-fno-math-errno", "-fPIC", "-w"]
VARIANTS = {"libsiril_ref.so": [], "libsiril_ref_fast.so": ["-fopenmp"]}


def load_hashes():
    h = {}
    if os.path.exists(HASHES_FILE):
        for line in open(HASHES_FILE):
            line = line.strip()
            if line and not line.startswith("#"):
                name, digest = line.split()
                h[name] = digest
    return h


def extract(update):
    os.makedirs(GEN, exist_ok=True)
    pinned = load_hashes()
    new = {}
    for name, (rel, a, b, _) in RANGES.items():
        path = os.path.join(REF, rel)
        with open(path, "r", encoding="utf-8") as f:
            lines = f.readlines()
        text = "".join(lines[a - 1:b])
        digest = hashlib.sha256(text.encode()).hexdigest()
        new[name] = digest
        if not update and name in pinned and pinned[name] != digest:
            raise SystemExit(f"reference range {name} ({rel}:{a}-{b}) changed: sha256 mismatch")
        with open(os.path.join(GEN, name + ".inc"), "w") as f:
            f.write(f"/* generated from {rel}:{a}-{b} (reference, not committed) */\n")
            f.write(text)
    if update or not pinned:
        with open(HASHES_FILE, "w") as f:
            f.write("# sha256 of the reference line ranges build_ref.py extracts (lock042/siril @ a5869d93)\n")
            for name in RANGES:
                f.write(f"{name} {new[name]}\n")
    # proto inlines are concatenated into one include
    with open(os.path.join(GEN, "proto_inlines.inc"), "w") as f:
        for name in RANGES:
            if name.startswith("proto_"):
                f.write(open(os.path.join(GEN, name + ".inc")).read() + "\n")


def run(cmd):
    print(" ".join(cmd))
    subprocess.check_call(cmd)


def build():
    inc = ["-I" + os.path.join(HERE, "shim"), "-I" + OUT, "-I" + os.path.join(REF, "src")]
    for so, extra in VARIANTS.items():
        odir = os.path.join(OUT, "obj_" + so.split(".")[0])
        os.makedirs(odir, exist_ok=True)
        objs = []
        for rel, cc in WHOLE_TUS:
            obj = os.path.join(odir, os.path.basename(rel).rsplit(".", 1)[0] + ".o")
            run([cc] + BASE + extra + inc + ["-c", os.path.join(REF, rel), "-o", obj])
            objs.append(obj)
        drv = os.path.join(odir, "ref_driver.o")
        run(["gcc"] + BASE + extra + inc + ["-c", os.path.join(HERE, "ref_driver.c"), "-o", drv])
        smp = os.path.join(odir, "ref_simple.o")
        run(["gcc"] + BASE + extra + inc + ["-c", os.path.join(HERE, "ref_simple.c"), "-o", smp])
        omp = os.path.join(odir, "ref_omp.o")
        run(["gcc"] + BASE + ["-fopenmp", "-c", os.path.join(HERE, "ref_omp.c"), "-o", omp])
        run(["g++", "-shared", "-fopenmp", "-o", os.path.join(OUT, so)] + objs + [drv, smp, omp, "-lm"])


def main():
    if not os.path.isdir(REF):
        print(f"{REF} not present: keeping prebuilt oracle/_ref as is")
        return 0
    extract("--update-hashes" in sys.argv)
    build()
    return 0


if __name__ == "__main__":
    sys.exit(main())
