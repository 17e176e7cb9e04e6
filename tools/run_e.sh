set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/e_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/e_pytest.log
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/e_ref.json 2> gpurun_out/e_ref.err
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/e_n1.json 2> gpurun_out/e_n1.err
timeout 300 python tools/probe.py --kind mixed --mib 1024 --levels 1,6 --reps 2 > gpurun_out/e_probe_mixed.log 2>&1
timeout 300 python tools/probe.py --kind fastq --mib 1024 --levels 1,6 --reps 2 --crc-mib 1024 > gpurun_out/e_probe_fastq.log 2>&1
tail -3 gpurun_out/e_pytest.log; cat gpurun_out/e_n1.json; cat gpurun_out/e_probe_mixed.log
