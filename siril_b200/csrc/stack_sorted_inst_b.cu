// explicit instantiations of the sorted engine kernel (split over files to build in parallel)
#include "stack_sorted.cuh"
SGPU_SORTED_VARIANT(3, 16)
SGPU_SORTED_VARIANT(3, 20)
SGPU_SORTED_VARIANT(3, 57)
SGPU_SORTED_VARIANT(4, 38)
SGPU_SORTED_VARIANT(5, 44)
