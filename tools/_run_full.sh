cd /root/repo
timeout 1500 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err
echo "rc=$?" >> gpurun_out/bench_default.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.json 2>> gpurun_out/bench_default.err
timeout 600 python bench.py --impl reference --workload finetune --steps 1 --warmup 1 > gpurun_out/bench_reference_ft.json 2>> gpurun_out/bench_default.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/md_launches.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-finetune > gpurun_out/md_ncu.log 2>&1
echo "ncu rc=$?" >> gpurun_out/bench_default.err
