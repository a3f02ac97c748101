/* TEST INFRASTRUCTURE ONLY — shim for src/core/proto.h: the seven rounding /
 * clamp inlines the hot path uses are extracted from the reference at build
 * time (gen/proto_inlines.inc); nothing else of proto.h is needed. */
#ifndef SHIM_PROTO_H
#define SHIM_PROTO_H
#include "core/siril.h"
#include "gen/proto_inlines.inc"
#endif
