#!/bin/bash
# usage: sass_hist.sh <obj> <mangled-name-substring>   -- opcode histogram of the biggest loop-free listing of a kernel
obj=$1; pat=$2
cuobjdump -sass "$obj" | awk -v pat="$pat" '/Function :/ {on = index($0, pat) > 0} on {print}' > /tmp/kernel.sass
grep -cE "^\s+/\*[0-9a-f]{4}\*/" /tmp/kernel.sass
grep -oE "^\s+/\*[0-9a-f]+\*/\s+(@!?U?P[0-9T]+ )?[A-Z0-9_.]+" /tmp/kernel.sass | awk '{print $NF}' | sed 's/\..*//' | sort | uniq -c | sort -rn | head -${3:-25}
