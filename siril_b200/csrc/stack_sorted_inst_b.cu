// explicit instantiations of the sorted engine kernel (split over files to build in parallel)
#include "stack_sorted.cuh"
SGPU_SORTED_VARIANT(3, 8, 4)
SGPU_SORTED_VARIANT(3, 25, 4)
SGPU_SORTED_VARIANT(3, 44, 4)
SGPU_SORTED_VARIANT(4, 13, 8)
SGPU_SORTED_VARIANT(4, 19, 4)
SGPU_SORTED_VARIANT(5, 44, 4)
