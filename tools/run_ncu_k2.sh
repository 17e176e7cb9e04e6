mkdir -p gpurun_out
for k in seg0 seg1; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_match -s 1 -c 1 -o gpurun_out/k2_$k -f python tools/probe.py --kind $k --mib 256 --levels 6 --reps 1 > gpurun_out/k2_$k.log 2>&1
  tail -2 gpurun_out/k2_$k.log
done
