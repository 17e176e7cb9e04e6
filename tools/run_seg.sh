mkdir -p gpurun_out
for k in seg0 seg1 seg2 seg3 seg4 seg5; do
  echo "== $k" >> gpurun_out/seg.log
  timeout 200 python tools/probe.py --kind $k --mib 256 --levels 6 --reps 1 >> gpurun_out/seg.log 2>&1
done
cat gpurun_out/seg.log
