#!/bin/bash
# Developer run (GPU box): the profile set of one build — bench lines of both arms, steady-state ncu launch
# list (4 steps after 100), ncu --set full of the k_pgsb launches of one step (details page + DRAM byte counters).
# usage: scripts/dev_profile_set.sh <tag>   -> gpurun_out/<tag>_*
tag=${1:-rXX}
out=gpurun_out
python bench.py > $out/${tag}_bench_n1.json 2> $out/${tag}_bench_n1.err
python bench.py --impl reference > $out/${tag}_bench_reference_n1.json 2> $out/${tag}_bench_reference_n1.err
STEPS=104 python scripts/dev_ncu_target.py > $out/${tag}_plain.log 2>&1 || exit 1
STEPS=104 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 2100 -c 84 --csv \
  --log-file $out/${tag}_launches_steady.csv python scripts/dev_ncu_target.py > $out/${tag}_ncu1.log 2>&1
STEPS=102 ncu --set full --clock-control none --import-source on -k regex:k_pgsb --launch-skip 800 --launch-count 8 \
  -o $out/${tag}_pgsb_full -f python scripts/dev_ncu_target.py > $out/${tag}_ncu2.log 2>&1
ncu -i $out/${tag}_pgsb_full.ncu-rep --page details > $out/${tag}_pgsb_full.txt 2>&1
ncu -i $out/${tag}_pgsb_full.ncu-rep --page raw --csv --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread,launch__grid_size,smsp__inst_executed.sum \
  > $out/${tag}_pgsb_raw.csv 2>&1
ls -la $out | tail -12
