#!/bin/bash
# A/B of two builds of the library on ONE box: B200_LIB_PATH selects the build, runs alternate
cd "$(dirname "$0")/.."
for rep in 1 2; do
  for v in old new; do
    if [ $v = old ]; then export B200_LIB_PATH=$PWD/sevennet_finetuning_b200/libold.so; else unset B200_LIB_PATH; fi
    python bench.py --no-alt --no-cpu-baseline --no-finetune --steps 6 --warmup 3 > gpurun_out/ab_${v}_${rep}.json 2> gpurun_out/ab_${v}_${rep}.err
    python - <<PY
import json
d = json.load(open("gpurun_out/ab_${v}_${rep}.json"))
k = d["kernel_ms_per_step"]
print("$v $rep", round(d["ms_per_step"], 2), {n: round(x, 2) for n, x in k.items() if x > 0.5})
PY
  done
done
