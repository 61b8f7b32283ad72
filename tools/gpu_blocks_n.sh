# spread of the PCG phases over the blocks of rank 0 at N ranks (C4 sharded)
N=${1:-2}
TSB_PCG_TIMING=1 TSB_PCG_TIMING_BLOCKS=1 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29771 tools/prof.py --workload c4 --lin 3 --pcg 20000 --gn 1 2>&1 | grep "linearize\|streamed PCG\|over .* blocks\|optimize\|P=" | cut -c1-330 | head -40
