// explicit instantiations of the selection engine kernel (split over files to build in parallel)
#include "stack_psel.cuh"
#define CE(a, b) LCE(a, b)
namespace sgpu {
SGPU_LOCALNET(8) {
#include "gen/sortnet_8.inc"
}
}  // namespace sgpu
SGPU_PSEL_VARIANT(4, 50, 37, 1)
SGPU_PSEL_VARIANT(4, 64, 50, 1)
SGPU_PSEL_VARIANT(4, 72, 64, 1)
