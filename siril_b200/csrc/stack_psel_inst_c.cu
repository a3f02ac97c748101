// explicit instantiation of the selection engine kernel (one file per group size: parallel builds)
#include "stack_psel.cuh"
#define CE(a, b) LCE(a, b)
namespace sgpu {
SGPU_LOCALNET(16) {
#include "gen/sortnet_16.inc"
}
SGPU_LOCALNET(24) {
#include "gen/sortnet_24.inc"
}
}  // namespace sgpu
SGPU_PSEL_VARIANT(2, 1)
