#!/bin/bash
# usage: tools/cfg_sweep.sh <lib> "<lanes warps level>" ...   ; env TASK, N pass through to tools/variant_probe.py
lib=$1; shift
export FLYB_LIB_PATH=$PWD/$lib
for cfg in "$@"; do
  set -- $cfg
  out=$(FLYB_LANES=$1 FLYB_WARPS=$2 FLYB_SMEM_LEVEL=$3 timeout 150 python tools/variant_probe.py 2>&1); rc=$?
  echo "cfg [$cfg] rc=$rc $(echo "$out" | tail -2 | tr '\n' ' ' | cut -c1-400)"
done
