// explicit instantiation of the selection engine kernel (one file per group size: parallel builds)
#include "stack_psel.cuh"
#define CE(a, b) LCE(a, b)
namespace sgpu {
SGPU_LOCALNET(4) {
#include "gen/sortnet_4.inc"
}
SGPU_LOCALNET(8) {
#include "gen/sortnet_8.inc"
}
}  // namespace sgpu
SGPU_PSEL_VARIANT(4, 1)
