// explicit instantiations of the lane-sort engine kernel (split over files to build in parallel)
#include "stack_lsort.cuh"
#define CE(a, b) LCE(a, b)
namespace sgpu {
SGPU_LOCALNET(20) {
#include "gen/sortnet_20.inc"
}
SGPU_LOCALNET(25) {
#include "gen/sortnet_25.inc"
}
SGPU_LOCALNET(32) {
#include "gen/sortnet_32.inc"
}
}  // namespace sgpu
SGPU_LSORT_VARIANT(2, 32, 25, 20, 4, 8)
SGPU_LSORT_VARIANT(2, 32, 32, 25, 4, 8)
SGPU_LSORT_VARIANT(2, 32, 25, 20, 2, 4)
SGPU_LSORT_VARIANT(2, 32, 25, 20, 4, 4)
SGPU_LSORT_VARIANT(2, 32, 25, 20, 2, 8)
