mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_inflate.py -m gpu -x -q 2>&1 | tail -25 | tee gpurun_out/inf_pytest.log
