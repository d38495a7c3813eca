#!/bin/bash
# A/B the per-kernel times of the headline step over several builds of the library (NMB_LIB_PATH):
# usage: tools/ab_libs.sh <lib.so> [<lib.so> ...]   ("default" = the in-tree build)
for so in "$@"; do
    if [ "$so" = default ]; then unset NMB_LIB_PATH; else export NMB_LIB_PATH=$(realpath $so); fi
    python bench.py --no-cpu-baseline --steps 100 --warmup 5 --sustained-s 0 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$so', d['ms_per_step'], d['roofline']['per_kernel_ms'])"
done
