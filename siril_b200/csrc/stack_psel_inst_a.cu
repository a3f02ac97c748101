// explicit instantiation of the selection engine kernel (one file per group size: parallel builds)
#include "stack_psel.cuh"
#define CE(a, b) LCE(a, b)
namespace sgpu {
SGPU_LOCALNET(64) {
#include "gen/sortnet_64.inc"
}
SGPU_LOCALNET(80) {
#include "gen/sortnet_80.inc"
}
}  // namespace sgpu
SGPU_PSEL_VARIANT(0, 4)
