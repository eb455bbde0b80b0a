#!/bin/bash
# usage: tools/sass_count.sh <obj-or-so>  -> SASS instruction count per function
cuobjdump -sass "$1" | grep -E "Function :|^\s+/\*[0-9a-f]{4,}\*/" | awk '/Function :/{name=$3; next} {cnt[name]++} END{for(n in cnt) print cnt[n], n}'
