#!/bin/bash
# Developer check (GPU box): bench.py under the given env assignments, one short line per configuration.
# usage: scripts/dev_bench_line.sh "GZCUDA_PGS_CHAIN_MIN=28" "GZCUDA_PGS_CHAIN_MIN=24 GZCUDA_PGS_SMEM_KB=44" ...
for cfg in "$@"; do
  env $cfg timeout 300 python bench.py --steps ${STEPS:-100} --warmup 20 2>/dev/null | tail -1 | \
    python -c "import sys,json; j=json.loads(sys.stdin.read()); print('%-60s value %.0f ms/step %.3f e2e %.0f' % ('$cfg', j['value'], j['ms_per_step'], j['e2e']['value']))"
done
