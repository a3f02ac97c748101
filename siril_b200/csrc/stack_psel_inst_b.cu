// explicit instantiations of the selection engine kernel (split over files to build in parallel)
#include "stack_psel.cuh"
#define CE(a, b) LCE(a, b)
namespace sgpu {
SGPU_LOCALNET(32) {
#include "gen/sortnet_32.inc"
}
}  // namespace sgpu
SGPU_PSEL_VARIANT(1, 50, 40, 2)
SGPU_PSEL_VARIANT(1, 64, 50, 2)
SGPU_PSEL_VARIANT(1, 80, 64, 2)
