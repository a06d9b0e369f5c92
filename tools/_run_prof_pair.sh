cd /root/repo
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-finetune"
$CMD > gpurun_out/pp_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:conv_pair_bwd_kernel -s 4 -c 4 -o gpurun_out/pp_bwd -f $CMD > gpurun_out/pp_ncu_bwd.log 2>&1
