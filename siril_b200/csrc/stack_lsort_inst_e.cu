// explicit instantiations of the lane-sort engine kernel (split over files to build in parallel)
#include "stack_lsort.cuh"
#define CE(a, b) LCE(a, b)
namespace sgpu {
SGPU_LOCALNET(20) {
#include "gen/sortnet_20.inc"
}
SGPU_LOCALNET(24) {
#include "gen/sortnet_24.inc"
}
}  // namespace sgpu
SGPU_LSORT_VARIANT(2, 20, 20, 16, 4, 8)
SGPU_LSORT_VARIANT(2, 24, 24, 20, 4, 8)
SGPU_LSORT_VARIANT(3, 20, 20, 16, 2, 8)
SGPU_LSORT_VARIANT(3, 24, 24, 20, 2, 8)
