mkdir -p gpurun_out
for b in 128 256 512 1024; do
  echo "== batch=$b" | tee -a gpurun_out/batch.log
  timeout 300 python tools/probe.py --kind mixed --mib 2048 --levels 6 --reps 2 --batch-mib $b 2>&1 | tee -a gpurun_out/batch.log
done
