#!/bin/bash
# A/B of experimental builds on ONE box: tools/ab_variants.sh <workload> <variant.so>... ; base first and last.
# Variants are built with PT_NVCC_EXTRA=... into gpupathtracer_b200/variants/ (git-ignored) and selected with PT_LIB_PATH.
wl=$1; shift
mkdir -p gpurun_out
run() { PT_LIB_PATH=$2 env $3 python bench.py --workload $wl --no-cpu-baseline --no-configs --steps ${STEPS:-64} --warmup 4 2>gpurun_out/ab_err.log | python -c "
import sys, json
for l in sys.stdin:
    l = l.strip()
    if l.startswith('{'):
        d = json.loads(l); print('$1', '$wl', round(d['value'], 1), d['unit'], 'ms/step', round(d['ms_per_step'], 4), 'e2e', round(d['e2e']['value'], 1), {k: round(v, 3) for k, v in (d.get('stage_ms_per_step') or {}).items()})
"; }
# arguments: variant .so files, or KEY=VALUE environment settings tried on the base build
run base ""
for v in "$@"; do
    case $v in
        *=*) run "$v" "" "$v" ;;
        *) run $(basename $v .so) $PWD/$v ;;
    esac
done
run base ""
