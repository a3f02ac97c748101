/* TEST INFRASTRUCTURE ONLY — shim for src/core/sequence_filtering.h */
#include "core/processing.h"
