// explicit instantiations of the sorted engine kernel (split over files to build in parallel)
#include "stack_sorted.cuh"
SGPU_SORTED_VARIANT(3, 32)
SGPU_SORTED_VARIANT(3, 38)
SGPU_SORTED_VARIANT(3, 44)
SGPU_SORTED_VARIANT(4, 50)
SGPU_SORTED_VARIANT(4, 57)
SGPU_SORTED_VARIANT(5, 57)
SGPU_SORTED_VARIANT(5, 64)
