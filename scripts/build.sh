#!/bin/bash
# builds libsgpu.so and says so only when the library is newer than every source
set -o pipefail
cd "$(dirname "$0")/../siril_b200/csrc" || exit 1
mkdir -p build; make -j8 > build/make.log 2>&1; rc=$?
grep -E "error|Error|Stop" build/make.log | head -20
if [ $rc -ne 0 ]; then echo "BUILD FAILED (rc=$rc)"; exit $rc; fi
newest=$(ls -t *.cu *.cuh ../../include/sgpu.h | head -1)
if [ "$newest" -nt ../libsgpu.so ]; then echo "BUILD STALE: $newest is newer than libsgpu.so"; exit 1; fi
echo "BUILD OK $(ls -la --time-style=+%H:%M:%S ../libsgpu.so | awk '{print $6}')"
