cd /root/repo
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-finetune --no-alt"
$CMD > gpurun_out/r02_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sectors.sum,lts__t_sectors_srcunit_tex_op_red.sum,smsp__inst_executed.sum --clock-control none -c 1200 --csv --log-file gpurun_out/r02_launches.csv $CMD > gpurun_out/r02_ncu1.log 2>&1
# the parity call of the bench (600 atoms) launches 36 conv_pair kernels before the first 100,000-atom step: skip them, then
# layer 0 (2 launches) -> the four forward launches of layer 1; the reverse pass starts 18 launches into a step with layer 4
timeout 900 ncu --set full --clock-control none --import-source on -k regex:conv_pair_fwd -s 20 -c 4 -o gpurun_out/r02_conv_pair_fwd -f $CMD > gpurun_out/r02_ncu2.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:conv_pair_bwd -s 22 -c 4 -o gpurun_out/r02_conv_pair_bwd -f $CMD > gpurun_out/r02_ncu3.log 2>&1
tail -n 2 gpurun_out/r02_ncu1.log; tail -n 2 gpurun_out/r02_ncu2.log; tail -n 2 gpurun_out/r02_ncu3.log
