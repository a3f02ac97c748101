// explicit instantiations of the lane-sort engine kernel (split over files to build in parallel)
#include "stack_lsort.cuh"
#define CE(a, b) LCE(a, b)
namespace sgpu {
SGPU_LOCALNET(19) {
#include "gen/sortnet_19.inc"
}
SGPU_LOCALNET(25) {
#include "gen/sortnet_25.inc"
}
SGPU_LOCALNET(32) {
#include "gen/sortnet_32.inc"
}
}  // namespace sgpu
SGPU_LSORT_VARIANT(4, 32, 25, 19, 1, 8)
SGPU_LSORT_VARIANT(4, 32, 32, 25, 1, 8)
SGPU_LSORT_VARIANT(5, 32, 32, 16, 1, 8)
