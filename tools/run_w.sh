mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_pipeline.py -m gpu -x -q > gpurun_out/w_pytest.log 2>&1; tail -3 gpurun_out/w_pytest.log
for w in 4 1 2 8; do
  if [ $w = 4 ]; then unset B2D_LIB; else export B2D_LIB=$PWD/tools/_bisect/libb2d_w$w.so; fi
  echo "== W=$w" | tee -a gpurun_out/w.log
  timeout 300 python tools/probe.py --kind mixed --mib 1024 --levels 1,6 --reps 2 2>&1 | tee -a gpurun_out/w.log
done
