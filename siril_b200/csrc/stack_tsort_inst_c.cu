// explicit instantiations of the thread-sort engine kernel (split over files to build in parallel)
#include "stack_tsort.cuh"
namespace sgpu {
SGPU_SORTNET(100) {
#include "gen/sortnet_100.inc"
}
}  // namespace sgpu
SGPU_TSORT_VARIANT(100, 3)
SGPU_TSORT_VARIANT(100, 2)
