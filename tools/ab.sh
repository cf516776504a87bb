#!/bin/bash
# same-box A/B of two builds of the library: ab/libA.so and ab/libB.so are swapped in turn under the in-tree name
# usage: tools/ab.sh '<command printing the numbers to compare>'
for round in 1 2 3; do
  for v in A B; do
    cp ab/lib$v.so fp8-emulation-toolkit_b200/libfpemu_b200.so
    echo "== $v round $round"
    bash -c "$1"
  done
done
