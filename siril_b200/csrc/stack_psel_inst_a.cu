// explicit instantiations of the selection engine kernel (split over files to build in parallel)
#include "stack_psel.cuh"
#define CE(a, b) LCE(a, b)
namespace sgpu {
SGPU_LOCALNET(50) {
#include "gen/sortnet_50.inc"
}
}  // namespace sgpu
SGPU_PSEL_VARIANT(0, 50, 40, 4)
SGPU_PSEL_VARIANT(0, 64, 50, 4)
SGPU_PSEL_VARIANT(0, 80, 64, 4)
