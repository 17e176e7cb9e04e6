# dev: parity of the pipeline against the CPU twin, then stage times per corpus kind
mkdir -p gpurun_out
T=${1:-ab}
timeout 600 python -m pytest tests/test_gpu_pipeline.py -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; tail -3 gpurun_out/${T}_pytest.log
timeout 300 python tools/probe.py --kind mixed --mib 1024 --levels 1,6 --reps 2 > gpurun_out/${T}_mixed.log 2>&1; cat gpurun_out/${T}_mixed.log
for k in seg0 seg1 seg2 seg3 seg4 seg5; do
  echo "== $k" >> gpurun_out/${T}_seg.log
  timeout 200 python tools/probe.py --kind $k --mib 256 --levels 6 --reps 1 >> gpurun_out/${T}_seg.log 2>&1
done
grep -E "==|stage" gpurun_out/${T}_seg.log
