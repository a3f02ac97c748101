// explicit instantiations of the selection engine kernel (split over files to build in parallel)
#include "stack_psel.cuh"
#define CE(a, b) LCE(a, b)
namespace sgpu {
SGPU_LOCALNET(16) {
#include "gen/sortnet_16.inc"
}
}  // namespace sgpu
SGPU_PSEL_VARIANT(3, 50, 40, 1)
SGPU_PSEL_VARIANT(3, 64, 50, 1)
SGPU_PSEL_VARIANT(3, 75, 64, 1)
