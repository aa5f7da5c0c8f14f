#!/bin/bash
# Developer run (GPU box): step time of the headline batch over the shared-memory classes of k_pgsb
for kb in "56 112" "64 112" "76 112" "76 200" "110 200"; do
  set -- $kb
  GZCUDA_PGSB_KB=$1 GZCUDA_PGSB_KB2=$2 python scripts/dev_step_time.py 2>&1 | head -1
done
