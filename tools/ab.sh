#!/bin/bash
# same-box A/B of two builds of the library: ab/libA.so and ab/libB.so are swapped in turn under the in-tree name
for round in 1 2; do
  for v in A B; do
    cp ab/lib$v.so fp8-emulation-toolkit_b200/libfpemu_b200.so
    echo "== $v round $round"
    python tools/bound16.py --only "$1" --reps 15 2>&1 | grep -E "$2" | cut -c1-120
    python tools/xchg_probe1.py --rounds 4 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read())['ms_per_sweep_median']; print({k: d[k] for k in ('plain','cast_plain_only','amax_only')})"
  done
done
