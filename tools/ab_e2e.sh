#!/bin/bash
# Dev tool: device-resident and end-to-end GB/s against batch size and number of compute lanes (A/B builds of libb2d
# under tools/_bisect, selected with B2D_LIB).  Usage on the GPU box: bash tools/ab_e2e.sh > gpurun_out/ab_e2e.log
for lib in "" tools/_bisect/libb2d_l3.so tools/_bisect/libb2d_l4.so; do
  for b in 32 64 128 256; do
    out=$(B2D_LIB=${lib:+$PWD/$lib} python bench.py --steps 2 --warmup 3 --no-extras --base-mib 64 --batch-mib $b 2>/dev/null)
    echo "lib=${lib:-default(2 lanes)} batch_mib=$b $(echo "$out" | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print("value %.2f e2e %.2f verify %s" % (d["value"], d["e2e"]["value"], d["verify_ok"]))')"
  done
done
