#!/bin/bash
# usage: tools/sass_hist.sh <kernel-name-substring>   -> opcode histogram of that kernel in libmobilert_b200.so
cuobjdump -sass mobileraytracer_b200/libmobilert_b200.so 2>/dev/null | awk -v pat="$1" '
/Function :/ { on = index($0, pat) > 0 }
on && /^\s+\/\*[0-9a-f]{4}\*\// { i=2; if ($2 ~ /^@/) i=3; op=$i; sub(/;/,"",op); split(op,a,"."); print a[1] }' | sort | uniq -c | sort -rn | head -${2:-30}
