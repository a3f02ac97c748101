// explicit instantiations of the sorted engine kernel (split over files to build in parallel)
#include "stack_sorted.cuh"
SGPU_SORTED_VARIANT(3, 4)
SGPU_SORTED_VARIANT(3, 8)
SGPU_SORTED_VARIANT(3, 13)
SGPU_SORTED_VARIANT(3, 64)
SGPU_SORTED_VARIANT(5, 38)
