#!/bin/bash
# usage: tools/bench_variants.sh tag "ENV1=.. ENV2=.." ...   (one short bench per variant, JSON lines under gpurun_out/)
mkdir -p gpurun_out
i=0
for v in "$@"; do
  i=$((i+1))
  env $v python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-extras --no-train > gpurun_out/variant_$i.json 2> gpurun_out/variant_$i.err
  echo "$v" > gpurun_out/variant_$i.env
done
