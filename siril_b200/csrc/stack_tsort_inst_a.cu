// explicit instantiations of the thread-sort engine kernel (split over files to build in parallel)
#include "stack_tsort.cuh"
namespace sgpu {
SGPU_SORTNET(8) {
#include "gen/sortnet_8.inc"
}
SGPU_SORTNET(16) {
#include "gen/sortnet_16.inc"
}
SGPU_SORTNET(24) {
#include "gen/sortnet_24.inc"
}
SGPU_SORTNET(32) {
#include "gen/sortnet_32.inc"
}
SGPU_SORTNET(50) {
#include "gen/sortnet_50.inc"
}
}  // namespace sgpu
SGPU_TSORT_VARIANT(8, 8)
SGPU_TSORT_VARIANT(16, 8)
SGPU_TSORT_VARIANT(24, 8)
SGPU_TSORT_VARIANT(32, 6)
SGPU_TSORT_VARIANT(50, 4)
