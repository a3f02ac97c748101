// explicit instantiations of the register engine kernel (split over files to build in parallel)
#include "stack_reg.cuh"
SGPU_REG_VARIANT(2, 64, 38, 4, false)
SGPU_REG_VARIANT(2, 64, 38, 4, true)
SGPU_REG_VARIANT(3, 64, 50, 2, false)
SGPU_REG_VARIANT(3, 64, 50, 2, true)
