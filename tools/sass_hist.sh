#!/bin/bash
# Opcode histogram of the hot kernels' SASS (cuobjdump), written to profiles/: tools/sass_hist.sh > profiles/r02_sass_opcodes.txt
cd "$(dirname "$0")/.."
for pat in "lk_level_phase_kernelILi7ELb1ELb1" "lk_level_phase1_kernelILb1ELb1" "pyr_down_fast_kernel"; do
  obj=optical-flow_b200/build/lk_level_phase.o
  [ "$pat" = "pyr_down_fast_kernel" ] && obj=optical-flow_b200/build/pyr_down.o
  echo "== $pat ($(basename $obj), sm_100a)"
  cuobjdump -sass $obj | awk -v pat="$pat" '/Function :/ {on = index($0, pat) > 0} on {print}' | grep -E '^\s+/\*[0-9a-f]{4}\*/' |
    sed -E 's/^\s+\/\*[0-9a-f]{4}\*\/\s+//; s/\s*\/\*.*$//; s/^@!?U?P[0-9T] +//' | awk '{print $1}' | sed 's/;$//' | sort | uniq -c | sort -rn |
    awk '{n+=$1; printf "%6d %s\n", $1, $2} END {printf "%6d total static instructions\n", n}'
done
