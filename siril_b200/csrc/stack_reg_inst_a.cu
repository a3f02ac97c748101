// explicit instantiations of the register engine kernel (split over files to build in parallel)
#include "stack_reg.cuh"
SGPU_REG_VARIANT(2, 25, 8, 6, false)
SGPU_REG_VARIANT(2, 25, 8, 6, true)
SGPU_REG_VARIANT(3, 38, 32, 3, false)
SGPU_REG_VARIANT(3, 38, 32, 3, true)
