// explicit instantiations of the thread-sort engine kernel (split over files to build in parallel)
#include "stack_tsort.cuh"
namespace sgpu {
SGPU_SORTNET(128) {
#include "gen/sortnet_128.inc"
}
}  // namespace sgpu
SGPU_TSORT_VARIANT(128, 2)
