cd /root/repo
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu -x 2>&1 | tail -4 > gpurun_out/t_all.log
timeout 900 python bench.py > gpurun_out/q_bench.json 2> gpurun_out/q_bench.err
timeout 600 python bench.py --workload finetune > gpurun_out/q_bench_ft.json 2> gpurun_out/q_bench_ft.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/q_launches.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-finetune > gpurun_out/q_ncu.log 2>&1
timeout 300 python tools/lin_time.py > gpurun_out/lin_time.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:irreps_linear_tc -s 2 -c 1 -o /tmp/lin -f python tools/lin_time.py > gpurun_out/ncu_lin.log 2>&1
ncu -i /tmp/lin.ncu-rep --page raw --csv > gpurun_out/q_lin_raw.csv 2>/dev/null
ncu -i /tmp/lin.ncu-rep --page source --csv > gpurun_out/q_lin_source.csv 2>/dev/null
