#!/bin/bash
# Dev tool: SASS of one kernel of a built library.  Usage: scripts/dev/sass.sh <lib.so> <kernel name substring>
L=$(cuobjdump -sass $1 2>/dev/null | grep -n "Function" | grep -A1 "$2" | cut -d: -f1 | tr '\n' ' '); set -- $1 $L
cuobjdump -sass $1 2>/dev/null | sed -n "$2,${3:-\$}p" | grep -E "/\*[0-9a-f]{4}\*/" | sed -E 's/^\s+//; s/\s+\/\* 0x[0-9a-f]+ \*\/\s*$//'
