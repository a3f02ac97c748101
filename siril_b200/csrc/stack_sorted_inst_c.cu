// explicit instantiations of the sorted engine kernel (split over files to build in parallel)
#include "stack_sorted.cuh"
SGPU_SORTED_VARIANT(3, 16, 4)
SGPU_SORTED_VARIANT(3, 32, 4)
SGPU_SORTED_VARIANT(3, 50, 4)
SGPU_SORTED_VARIANT(4, 38, 4)
SGPU_SORTED_VARIANT(4, 50, 4)
SGPU_SORTED_VARIANT(5, 50, 4)
SGPU_SORTED_VARIANT(4, 7, 8)
