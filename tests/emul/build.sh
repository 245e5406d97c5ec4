#!/bin/sh
# Builds the test-only emulation library (see cuda_emul.h).  Usage: tests/emul/build.sh
set -e
cd "$(dirname "$0")/../.."
mkdir -p tests/emul/_build
FLAGS="-std=c++20 -O1 -g -pthread -fPIC -shared -Wall -Wno-unused-function -Wno-unknown-pragmas -ffp-contract=off"
# stress variant: tiny subs and no run-in, so that nearly every sub needs the Jacobi repair path
g++ $FLAGS -DB2_X_FORCE_SHIFT=9 -o tests/emul/_build/libb200jpeg_emul_stress.so \
    tests/emul/emul_lib.cpp jpeg_decoder_b200/csrc/host_parse.cpp
g++ -std=c++20 -O1 -g -pthread -fPIC -shared -Wall -Wno-unused-function -Wno-unknown-pragmas -ffp-contract=off \
    -o tests/emul/_build/libb200jpeg_emul.so tests/emul/emul_lib.cpp jpeg_decoder_b200/csrc/host_parse.cpp
