cd /root/repo
timeout 900 ncu --profile-from-start off --set full --import-source on --clock-control none -k regex:"irreps_linear_tc_kernel|wgrad_splitk" --launch-skip 20 --launch-count 24 -o /tmp/ft_lin_full python tools/ft_ncu_target.py > gpurun_out/ft_ncu_full.log 2>&1
echo "ncu rc=$?" > gpurun_out/ft_ncu.err
ncu -i /tmp/ft_lin_full.ncu-rep --page raw --csv > gpurun_out/ft_lin_full_raw.csv 2>/dev/null
ls -la /tmp/ft_lin_full.ncu-rep >> gpurun_out/ft_ncu.err
