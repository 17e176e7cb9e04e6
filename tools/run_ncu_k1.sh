mkdir -p gpurun_out
for w in 4 1; do
  if [ $w = 4 ]; then unset B2D_LIB; else export B2D_LIB=$PWD/tools/_bisect/libb2d_w$w.so; fi
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_chains -s 1 -c 1 -o gpurun_out/k1_w$w -f python tools/probe.py --kind mixed --mib 256 --levels 6 --reps 1 > gpurun_out/k1_w$w.log 2>&1
  tail -2 gpurun_out/k1_w$w.log
done
ls -la gpurun_out
