# usage: bash tools/_run_n.sh N  (under gpurun --gpus N)
cd /root/repo
mkdir -p gpurun_out
N=$1
python -m pytest tests/test_gpu_parallel.py -x -q -m gpu -k nccl -s 2>&1 | tail -6 > gpurun_out/n${N}_nccl_test.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29555 tools/check_parallel.py 20000 > gpurun_out/n${N}_check_parallel.log 2>&1
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29556 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/n${N}_bench.json 2> gpurun_out/n${N}_bench.err
tail -4 gpurun_out/n${N}_nccl_test.log; grep "rank" gpurun_out/n${N}_check_parallel.log | tail -8; tail -2 gpurun_out/n${N}_bench.err
