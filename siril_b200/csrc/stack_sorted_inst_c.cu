// explicit instantiations of the sorted engine kernel (split over files to build in parallel)
#include "stack_sorted.cuh"
SGPU_SORTED_VARIANT(3, 25)
SGPU_SORTED_VARIANT(3, 50)
SGPU_SORTED_VARIANT(4, 44)
SGPU_SORTED_VARIANT(4, 64)
SGPU_SORTED_VARIANT(5, 50)
