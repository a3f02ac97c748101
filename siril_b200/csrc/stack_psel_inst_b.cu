// explicit instantiation of the selection engine kernel (one file per group size: parallel builds)
#include "stack_psel.cuh"
#define CE(a, b) LCE(a, b)
namespace sgpu {
SGPU_LOCALNET(32) {
#include "gen/sortnet_32.inc"
}
SGPU_LOCALNET(48) {
#include "gen/sortnet_48.inc"
}
}  // namespace sgpu
SGPU_PSEL_VARIANT(1, 2)
