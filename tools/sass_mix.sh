#!/bin/bash
# usage: tools/sass_mix.sh <object> <kernel-substring>  -- opcode histogram of one kernel
cuobjdump -sass "$1" 2>/dev/null | awk -v k="$2" '/Function : /{f=(index($0,k)>0)} f' | grep -E "^\s+/\*[0-9a-f]{4}\*/" | sed -E 's/^\s+\/\*[0-9a-f]+\*\/\s+//' | awk '{op=$1; if (op ~ /^@/) op=$2; sub(/\..*/,"",op); print op}' | sort | uniq -c | sort -rn | head -${3:-25}
