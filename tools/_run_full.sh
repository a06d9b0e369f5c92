cd /root/repo
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -x 2>&1 | tail -4 > gpurun_out/t_all.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1
timeout 300 python bench.py --workload finetune --no-cpu-baseline > gpurun_out/ft.json 2> gpurun_out/ft.err
