timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log | cut -c1-200
for w in c3 c4 c5; do echo "--- $w phased (B streamed)"; TSB_PCG_PHASED=1 timeout 600 python tools/prof.py --workload $w --lin 1 --pcg 32 2>&1 | tail -1 | cut -c1-160; done
echo "--- c4 phased unit"; TSB_NO_STREAM=1 TSB_PCG_PHASED=1 timeout 600 python tools/prof.py --workload c4 --lin 1 --pcg 32 2>&1 | tail -1 | cut -c1-160
echo "--- mgpu"; timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/mgpu_check.py 2>&1 | grep "world\|oracle\|replicated\|MGPU" | tail -5
echo "--- c3 2 ranks sharded, 3 GN iterations"; timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29514 tools/prof.py --workload c3 --lin 1 --pcg 20000 --gn 3 2>&1 | grep "optimize" | cut -c1-200
echo "--- c3 1 rank, 3 GN iterations"; timeout 900 python tools/prof.py --workload c3 --lin 1 --pcg 20000 --gn 3 2>&1 | grep "optimize" | cut -c1-200
echo "--- c3 1 rank no stream, 3 GN iterations"; TSB_NO_STREAM=1 timeout 900 python tools/prof.py --workload c3 --lin 1 --pcg 20000 --gn 3 2>&1 | grep "optimize" | cut -c1-200
