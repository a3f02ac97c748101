// explicit instantiations of the selection engine kernel (split over files to build in parallel)
#include "stack_psel.cuh"
#define CE(a, b) LCE(a, b)
namespace sgpu {
SGPU_LOCALNET(24) {
#include "gen/sortnet_24.inc"
}
}  // namespace sgpu
SGPU_PSEL_VARIANT(2, 50, 40, 1)
SGPU_PSEL_VARIANT(2, 64, 50, 1)
