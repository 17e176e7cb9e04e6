# last check of a commit on one B200: GPU tests, smoke(), default bench
mkdir -p gpurun_out
T=${1:-r1h}
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest.log
tail -3 gpurun_out/${T}_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 600 python bench.py > gpurun_out/bench_${T}_n1.json 2> gpurun_out/${T}_n1.err
cut -c1-200 gpurun_out/bench_${T}_n1.json
