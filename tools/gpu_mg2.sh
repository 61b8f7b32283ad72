# round 2: sharded path with the NVLink peer exchange on N GPUs (N = $1, default 2)
N=${1:-2}
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader | head -1
for args in "--workload c2 --scale 0.2" "--workload c3 --scale 0.1 --iters 3" "--workload c5 --scale 0.02 --iters 3"; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29751 tools/mgpu_check.py $args 2>&1 | grep -v "^\s*$\|OMP_NUM\|\*\*\*\*" | tail -6
done
echo "--- c3 full, $N ranks"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29752 tools/prof.py --workload c3 --lin 2 --pcg 20000 --gn 3 2>&1 | grep "optimize\|linearize\|solve" | head -12
