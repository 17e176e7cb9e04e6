mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_inflate.py -m gpu -x -q 2>&1 | tail -3 | tee gpurun_out/inf_pytest.log
timeout 300 python tools/probe.py --kind mixed --mib 1024 --levels 6 --reps 2 --inflate 2>&1 | grep -E "inflate" | tee gpurun_out/inf_probe.log
timeout 300 python tools/probe.py --kind mixed --mib 4096 --levels 6 --reps 1 --inflate 2>&1 | grep -E "inflate" | tee -a gpurun_out/inf_probe.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:k_inf --csv --log-file gpurun_out/inf_launches.csv python tools/probe.py --kind mixed --mib 4096 --levels 6 --reps 1 --inflate > /dev/null 2>&1
grep k_inf gpurun_out/inf_launches.csv | awk -F'","' '{split($5,a,"("); print a[1], $NF}'
