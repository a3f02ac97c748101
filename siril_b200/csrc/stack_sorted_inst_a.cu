// explicit instantiations of the sorted engine kernel (split over files to build in parallel)
#include "stack_sorted.cuh"
SGPU_SORTED_VARIANT(3, 4, 4)
SGPU_SORTED_VARIANT(3, 13, 4)
SGPU_SORTED_VARIANT(3, 38, 4)
SGPU_SORTED_VARIANT(4, 19, 8)
SGPU_SORTED_VARIANT(5, 38, 4)
