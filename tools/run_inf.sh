mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_inflate.py -m gpu -x -q 2>&1 | tail -3 | tee gpurun_out/inf_pytest.log
timeout 120 python tools/probe.py --kind mixed --mib 256 --levels 6 --reps 2 --inflate 2>&1 | grep -E "inflate" | tee gpurun_out/inf_probe.log
timeout 200 python tools/probe.py --kind mixed --mib 4096 --levels 6 --reps 2 --inflate 2>&1 | grep -E "inflate" | tee -a gpurun_out/inf_probe.log
