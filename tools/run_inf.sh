mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_inflate.py tests/test_gpu_api.py tests/test_gpu_pipeline.py -m gpu -x -q 2>&1 | tail -15 | tee gpurun_out/inf_pytest.log
