"""siril_b200 — B200-native (sm_100a) rejection-stacking hot path of Siril behind a C ABI.

The product is `libsgpu.so` (CUDA kernels + C ABI, include/sgpu.h).  This package is the thin
Python host mirror of the reference's interface for that path (struct stacking_args, the
rejection / normalization / weighting enums, stack_mean_with_rejection / stack_median /
do_normalization), used by the tests and the benchmark.  There is no CPU fallback: importing
works anywhere, computing needs the CUDA library and a GPU and fails loudly otherwise.
"""
from .enums import *  # noqa: F401,F403
from .stacking import (StackingArgs, NormCoeff, main_stack, stack_mean_with_rejection, stack_median,
                       stack_summing_generic, stack_addmax, stack_addmin,  # noqa: F401
                       do_normalization, StackResult, gesdt_critical_values, result_bgnoise)
from ._lib import lib, SgpuError, Context  # noqa: F401
