"""Per-op, per-shape CUDA-event times of one fine-tuning iteration (config 2 batch)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import argparse, torch
import bench
from sevennet_finetuning_b200 import ops

a = argparse.Namespace(batch=32, steps=2)
ops._TIMING_SHAPES = True
orig = ops.timing_summary
res = {}
def patched():
    r = orig()
    res.update(r)
    return r
ops.timing_summary = patched
line = bench.finetune_bench(a, 0, 1, 0, 3, cpu_baseline=False, clocks=False)
tot = 0
for k, (ms, calls) in sorted(res.items(), key=lambda kv: -kv[1][0]):
    print(f'{ms / 2:8.3f} ms  {calls // 2:4d} calls  {k}')
    tot += ms / 2
print('sum', tot, 'step', line['ms_per_step'])
