#!/bin/bash
# tools/sweep_knobs.sh [workload]: run-time knob sweep of the default bench on ONE box (no rebuild); base first and last.
wl=${1:-atrium}
mkdir -p gpurun_out
run() { env "$@" python bench.py --workload $wl --no-cpu-baseline --steps ${STEPS:-64} --warmup 4 2>>gpurun_out/sweep_err.log | python -c "
import sys, json
for l in sys.stdin:
    l = l.strip()
    if l.startswith('{'):
        d = json.loads(l); print('$*', round(d['value'], 1), 'ms/step', round(d['ms_per_step'], 4), 'e2e', round(d['e2e']['value'], 1))
"; }
run PT_X=base
run PT_FRAME_BLOCKS_PER_SM=3
run PT_FRAME_BLOCKS_PER_SM=5
run PT_FRAME_BLOCKS_PER_SM=6
run PT_SLICES=2
run PT_SLICES=4
run PT_SLICES=4 PT_FRAME_BLOCKS_PER_SM=3
run PT_INNER_BURST=8
run PT_INNER_BURST=16
run PT_LEAF_BURST=3
run PT_LEAF_THRESH=6
run PT_FETCH_THRESH=8
run PT_FETCH_THRESH=16
run PT_SHADE_BLOCKS_PER_SM=4
run PT_X=base
