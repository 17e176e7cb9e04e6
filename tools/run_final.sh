# round-end evidence on one B200: GPU tests, both bench arms, ncu launch list (after the plain runs)
mkdir -p gpurun_out
T=${1:-r1e}
timeout 1200 python -m pytest tests -m gpu -x -q --durations=8 > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest.log
tail -14 gpurun_out/${T}_pytest.log
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_${T}_reference.json 2> gpurun_out/${T}_ref.err
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_${T}_n1.json 2> gpurun_out/${T}_n1.err
cat gpurun_out/bench_${T}_n1.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_${T}.csv python bench.py --steps 2 --warmup 1 --no-extras > gpurun_out/${T}_ncu_bench.log 2>&1
tail -2 gpurun_out/launches_${T}.csv
B2D_INF_SERIAL=1 timeout 600 ncu --set full --import-source on --clock-control none -k regex:k_inf -c 3 -o gpurun_out/prof_${T}_inf -f python tools/probe.py --kind mixed --mib 4096 --levels 6 --reps 1 --inflate > gpurun_out/${T}_ncu_inf.log 2>&1
tail -3 gpurun_out/${T}_ncu_inf.log
