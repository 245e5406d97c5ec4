#!/bin/bash
# Dev tool: build several -D variants in parallel.  Usage: scripts/dev/build_variants.sh "name flags..." "name2 flags..." ...
rm -rf build/variants; mkdir -p build/variants
for v in "$@"; do set -- $v; n=$1; shift; bash scripts/build_variant.sh $n "$@" 2>&1 | grep -v "warning\|\^\|^$\|Remark\|slot_a" & done; wait; ls build/variants
