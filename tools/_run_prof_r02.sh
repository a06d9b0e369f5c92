cd /root/repo
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-finetune --no-alt"
$CMD > gpurun_out/r02_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sectors.sum,lts__t_sectors_srcunit_tex_op_red.sum,smsp__inst_executed.sum --clock-control none -c 1200 --csv --log-file gpurun_out/r02_launches.csv $CMD > gpurun_out/r02_ncu1.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:conv_pair -s 12 -c 12 -o gpurun_out/r02_conv_pair -f $CMD > gpurun_out/r02_ncu2.log 2>&1
tail -2 gpurun_out/r02_ncu1.log gpurun_out/r02_ncu2.log
