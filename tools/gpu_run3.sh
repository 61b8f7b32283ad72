set -x
nvidia-smi -L
timeout 600 python -m pytest tests -m gpu -x -q -k "two_ranks" > gpurun_out/pytest_2rank.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_2rank.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/mgpu_check.py --workload c2 --scale 0.2 > gpurun_out/mgpu_check_2.log 2>&1; echo "mgpu rc=$?"; tail -4 gpurun_out/mgpu_check_2.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 2 --warmup 1 --workload c2 > gpurun_out/bench_c2_n2.log 2> gpurun_out/bench_c2_n2.err; echo "bench c2 n2 rc=$?"; cat gpurun_out/bench_c2_n2.log; tail -3 gpurun_out/bench_c2_n2.err
