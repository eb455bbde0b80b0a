#!/bin/bash
# usage: tools/sass_layout.sh <obj> : start address of every subroutine inside the step kernel's text
d=$(mktemp -d); (cd $d && cuobjdump -xelf all "$(realpath $1)" >/dev/null && nvdisasm -c *.cubin > dis.txt)
awk '/^\.text\._ZN2bb14bb_step_kernel/{f=1} f' $d/dis.txt | grep -E "^\\\$|^\s+/\*[0-9a-f]{4,}\*/" | awk '/^\$/{print last, $0; next} {last=$1}'
awk '/^\.text\._ZN2bb14bb_step_kernel/{f=1} f' $d/dis.txt | grep -E "^\s+/\*[0-9a-f]{4,}\*/" | tail -1 | cut -c1-40
