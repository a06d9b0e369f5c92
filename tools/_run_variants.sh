cd /root/repo
mkdir -p gpurun_out
: > gpurun_out/variants.log
for so in sevennet_finetuning_b200/libsevennet_b200*.so; do
  B200_LIB_PATH=$PWD/$so timeout 300 python bench.py --steps 5 --warmup 2 --no-cpu-baseline --no-finetune > gpurun_out/v.json 2> gpurun_out/v.err
  python - "$so" <<'PY' >> gpurun_out/variants.log
import json,sys
try:
    d=json.loads(open('gpurun_out/v.json').read().strip().splitlines()[-1])
    k=d['kernel_ms_per_step']
    print(sys.argv[1].split('/')[-1], 'step %.2f fwd %.2f bwd %.2f node %.2f' % (d['ms_per_step'], k['conv_fwd'], k['conv_bwd_edge'], k['node_linear_gate']))
except Exception as e:
    print(sys.argv[1], 'FAILED', e, open('gpurun_out/v.err').read()[-300:])
PY
done
cat gpurun_out/variants.log
