/* TEST HARNESS ONLY — src/core/proto.h: the rounding inlines, extracted from the reference at build time */
#ifndef HARNESS_PROTO_H
#define HARNESS_PROTO_H
#include "core/siril.h"
#include "gen/proto_inlines.inc"
#endif
