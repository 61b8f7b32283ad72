# sharded path on N GPUs: N-vs-1 (and oracle) parity on small graphs, then the bench line
N=${1:-4}
for args in "--workload c2 --scale 0.2" "--workload c3 --scale 0.1 --iters 3" "--workload c5 --scale 0.02 --iters 3 --oracle 0 --state-tol 1e-7"; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29751 tools/mgpu_check.py $args 2>&1 | grep -v "^\s*$\|OMP_NUM\|\*\*\*\*" | tail -14 | cut -c1-300
done
bash tools/gpu_bench2.sh $N
