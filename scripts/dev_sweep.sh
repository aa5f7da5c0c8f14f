#!/bin/bash
# usage: dev_sweep.sh "SMEM_KB=44,RT=28 SMEM_KB=100 ..."   (each item: comma-separated GZCUDA_PGS_* knobs)
for cfg in $1; do
  envs=""
  for kv in ${cfg//,/ }; do envs="$envs GZCUDA_PGS_$kv"; done
  env $envs python scripts/dev_pgs_knobs.py 2>&1 | tail -1
done
