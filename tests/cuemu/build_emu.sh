#!/bin/sh
# TEST/DEBUG HARNESS ONLY: compiles the kernel sources for the host-thread emulator (see cuemu.h).
# Skipped when the library is newer than every source it is built from (three test modules ask for it).
set -e
HERE=$(cd "$(dirname "$0")" && pwd)
SRC=$HERE/../../xlayer_prover_b200/csrc
OUT=$HERE/libglprover_emu.so
if [ -f "$OUT" ] && [ -z "$(find "$SRC" "$HERE" "$HERE/../../include" \( -name '*.cu' -o -name '*.cuh' -o -name '*.h' -o -name '*.cpp' -o -name '*.sh' \) -newer "$OUT" 2>/dev/null)" ]; then
    echo "up to date: $OUT"
    exit 0
fi
/usr/bin/g++ -std=c++17 -O2 -g -fPIC -shared -pthread -DGLP_EMU -Wall -Wno-unused-function -Wno-unknown-pragmas \
    -I"$HERE" -I"$SRC" -x c++ "$SRC/capi.cu" "$HERE/cuemu.cpp" -o "$OUT.tmp.$$"
mv "$OUT.tmp.$$" "$OUT"
echo "built $OUT"
