# one ncu --set full capture of every pipeline kernel on one 256 MiB batch of the bench workload (level 6)
mkdir -p gpurun_out
T=${1:-r1e}
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_chains|k_match|k_parse|k_plan|k_scan|k_emit' -s 6 -c 6 -o gpurun_out/prof_$T -f python tools/probe.py --kind mixed --mib 256 --levels 6 --reps 1 > gpurun_out/prof_$T.log 2>&1
tail -3 gpurun_out/prof_$T.log
