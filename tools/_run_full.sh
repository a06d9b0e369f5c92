cd /root/repo
timeout 1200 ncu --set full --import-source on --clock-control none -k regex:"conv_bwde_kernel|conv_fwd_kernel" --launch-skip 26 --launch-count 26 -o /tmp/conv_tab python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-finetune > gpurun_out/conv_ncu.log 2>&1
echo "ncu rc=$?" > gpurun_out/conv_ncu.err
ncu -i /tmp/conv_tab.ncu-rep --page raw --csv > gpurun_out/conv_tab_raw.csv 2>/dev/null
