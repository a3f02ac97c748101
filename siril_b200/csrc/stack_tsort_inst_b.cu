// explicit instantiations of the thread-sort engine kernel (split over files to build in parallel)
#include "stack_tsort.cuh"
namespace sgpu {
SGPU_SORTNET(64) {
#include "gen/sortnet_64.inc"
}
SGPU_SORTNET(80) {
#include "gen/sortnet_80.inc"
}
}  // namespace sgpu
SGPU_TSORT_VARIANT(64, 4)
SGPU_TSORT_VARIANT(80, 3)
