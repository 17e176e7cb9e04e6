# 8 GPUs of one box: the scaling line of bench.py and BASELINE.json configs[4] (16 GiB streaming write, end to end)
mkdir -p gpurun_out
N=${1:-8}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_r1g_n$N.json 2> gpurun_out/r1g_n$N.err
cat gpurun_out/bench_r1g_n$N.json
DEV=$(python -c "print(','.join(str(i) for i in range($N)))")
nproc > gpurun_out/stream_r1g_n$N.log
B2D_DEVICES=$DEV timeout 600 python tools/stream_bench.py --gib 16 --threads $(nproc) --reps 3 >> gpurun_out/stream_r1g_n$N.log 2>&1
echo "-- reference, same host, threads=$(nproc), 2 GiB sample" >> gpurun_out/stream_r1g_n$N.log
timeout 300 python tools/stream_bench.py --ref --gib 2 --threads $(nproc) --reps 1 >> gpurun_out/stream_r1g_n$N.log 2>&1
cat gpurun_out/stream_r1g_n$N.log
