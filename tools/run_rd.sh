for w in 256 512 1024; do echo "== window $w MiB"; B2D_READ_WINDOW_MIB=$w timeout 300 python tools/stream_bench.py --read --gib 8 --threads 16 --reps 3 2>&1 | grep -E "rep2|rep1"; done
