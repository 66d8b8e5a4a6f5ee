#!/bin/sh
# TEST/DEBUG HARNESS ONLY: compiles the kernel sources for the host-thread emulator (see cuemu.h).
set -e
HERE=$(cd "$(dirname "$0")" && pwd)
SRC=$HERE/../../xlayer_prover_b200/csrc
/usr/bin/g++ -std=c++17 -O2 -g -fPIC -shared -pthread -DGLP_EMU -Wall -Wno-unused-function -Wno-unknown-pragmas \
    -I"$HERE" -I"$SRC" -x c++ "$SRC/capi.cu" "$HERE/cuemu.cpp" -o "$HERE/libglprover_emu.so"
echo "built $HERE/libglprover_emu.so"
