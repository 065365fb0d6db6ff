#!/bin/bash
# A/B helper for GPU runs: tests/ab.sh SIZE variant[@PF[@NW]] ...  (variant = abtest/<name>.so,
# PF = DES_PF_ELEMS, NW = DES_NODE_WINDOW for that run).  Each variant is copied over the
# in-tree library.
size=$1; shift
mkdir -p gpurun_out
for spec in "$@"; do
  IFS=@ read -r v pf nw <<< "$spec"
  cp abtest/$v.so dynearthsol_b200/csrc/libdes_b200.so
  tag=${size}_${v}_pf${pf:-def}_nw${nw:-def}
  env ${pf:+DES_PF_ELEMS=$pf} ${nw:+DES_NODE_WINDOW=$nw} timeout 600 python bench.py --size $size --steps 30 --warmup 3 --no-cpu --no-e2e > gpurun_out/ab_$tag.json 2> gpurun_out/ab_$tag.err
  python tests/bench_summary.py gpurun_out/ab_$tag.json
done
