#!/bin/bash
# device SAH builder variants (bins per axis, small-subtree threshold): build time, SAH cost, frame rate / primary-ray rate
for v in "" "$@"; do
  echo "== ${v:-default}"
  PT_LIB_PATH=${v:+$PWD/$v} EXP_ONLY=sah,build python tools/exp_builders.py atrium 2>&1 | grep "device SAH\|host SBVH"
  PT_LIB_PATH=${v:+$PWD/$v} EXP_ONLY=sah,build python tools/exp_builders_bunny.py 2>&1 | grep "device SAH\|host SBVH"
done
