#!/bin/bash
# Round-2 profile captures (run under gpurun, one GPU).  Every capture follows a plain run of the same command that exited 0.
set -u
OUT=gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-grids"
$CMD > $OUT/r2_bench_for_ncu.json 2> $OUT/r2_bench_for_ncu.err || exit 1
# (1) launch list of the default bench
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $OUT/r2_launches_bench_default.csv $CMD > $OUT/ncu1.log 2>&1
# (2) one steady-state LM iteration at 128^3, full sets
ncu --set full --clock-control none --import-source on -k regex:"k_sweep|k_dist|k_small" -s 16 -c 8 -o $OUT/r2_iter_128 $CMD > $OUT/ncu2.log 2>&1
ncu -i $OUT/r2_iter_128.ncu-rep --page raw --csv > $OUT/r2_iter_128_raw.csv 2>/dev/null
# (3) 512^3: the sweeps and the general distance solve
CMD5="python bench.py --steps 2 --warmup 2 --no-cpu-baseline --no-grids --grid 512"
$CMD5 > $OUT/r2_bench512_for_ncu.json 2> $OUT/r2_bench512_for_ncu.err || exit 1
ncu --set full --clock-control none --import-source on -k regex:"k_sweep|k_dist_pcg|k_small" -s 16 -c 8 -o $OUT/r2_iter_512 $CMD5 > $OUT/ncu3.log 2>&1
ncu -i $OUT/r2_iter_512.ncu-rep --page raw --csv > $OUT/r2_iter_512_raw.csv 2>/dev/null
ls -la $OUT/*.csv | tail -5
