#!/bin/bash
# usage: tools/spill_lines.sh <object file> <kernel name pattern>   -- source lines of the local-memory (spill) accesses of a kernel
obj=$(readlink -f "$1"); tmp=$(mktemp -d); cd $tmp || exit 1
cuobjdump -xelf all "$obj" > /dev/null || exit 1
nvdisasm -g -c ./*.cubin > all.dis
awk -v pat="$2" '/\.section[ \t]+\.text\./{f = ($0 ~ pat)} f' all.dis > k.dis
echo "instructions: $(grep -c '^\s*/\*[0-9a-f]\{4\}\*/' k.dis)"
grep -n "LDL\|STL\|//## File" k.dis | grep -B1 "LDL\|STL" | grep -v "^--" | sed 's/.*csrc\///' 
rm -rf $tmp
