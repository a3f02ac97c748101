// explicit instantiations of the lane-sort engine kernel (split over files to build in parallel)
#include "stack_lsort.cuh"
#define CE(a, b) LCE(a, b)
namespace sgpu {
SGPU_LOCALNET(8) {
#include "gen/sortnet_8.inc"
}
SGPU_LOCALNET(13) {
#include "gen/sortnet_13.inc"
}
SGPU_LOCALNET(16) {
#include "gen/sortnet_16.inc"
}
}  // namespace sgpu
SGPU_LSORT_VARIANT(1, 8, 8, 0, 4, 8)
SGPU_LSORT_VARIANT(2, 8, 8, 4, 4, 8)
SGPU_LSORT_VARIANT(2, 16, 13, 8, 4, 8)
SGPU_LSORT_VARIANT(2, 16, 16, 13, 4, 8)
