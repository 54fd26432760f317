#!/bin/sh
# tests/simt_emu/build.sh -- build the host emulation of the device code (TEST HARNESS ONLY)
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
g++ -O2 -g -std=c++17 -fPIC -shared $RMPC_EMU_FLAGS -Wall -Wno-unused-variable -Wno-unused-function \
    -I"$HERE" -I"$HERE/../../robotic-mpc_b200/csrc" "$HERE/emu_driver.cpp" "$HERE/simt_emu.cpp" -o "$HERE/librmpc_emu.so"
