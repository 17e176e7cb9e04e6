#!/bin/bash
# Round-2 profiler pass (run on the GPU box AFTER the plain bench exited 0; numbers printed under ncu are never bench values):
#   1. launch list of the bench command             -> gpurun_out/launches_r2.csv
#   2. ncu --set full of one launch of every deflate kernel on a 256 MiB level-6 batch (and the inflate kernels)
#                                                   -> gpurun_out/prof_r2.ncu-rep + raw csv
set -x
python bench.py --steps 2 --warmup 3 --no-extras > gpurun_out/ncu_plain_bench.json 2> gpurun_out/ncu_plain_bench.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_r2.csv \
    python bench.py --steps 2 --warmup 3 --no-extras > gpurun_out/ncu_launch.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'k_chains|k_match|k_parse|k_plan|k_scan|k_emit|k_inf' -s 66 -c 12 \
    -o gpurun_out/prof_r2 -f python tools/ncu_target.py > gpurun_out/ncu_full.log 2>&1
ncu -i gpurun_out/prof_r2.ncu-rep --page raw --csv > gpurun_out/prof_r2_raw.csv 2> /dev/null
