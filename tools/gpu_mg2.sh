# round 2: sharded path with the NVLink peer exchange on 2 GPUs
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader
nvidia-smi topo -m | head -6
for args in "--workload c2 --scale 0.2" "--workload c3 --scale 0.1 --iters 3"; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29751 tools/mgpu_check.py $args 2>&1 | grep -v "^\s*$" | tail -8
done
