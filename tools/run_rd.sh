mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_inflate.py tests/test_gpu_api.py -m gpu -x -q 2>&1 | tail -4
timeout 300 python tools/stream_bench.py --read --gib 4 --threads 16 --reps 4 2>&1 | tee gpurun_out/read.log
