timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
echo "--- custom barrier"; timeout 300 python tools/prof.py --workload c2 --lin 1 --pcg 4000 --gn 50 2>&1 | tail -1 | cut -c1-300
echo "--- cg grid.sync"; TSB_RES_CGSYNC=1 timeout 300 python tools/prof.py --workload c2 --lin 1 --pcg 4000 --gn 50 2>&1 | tail -1 | cut -c1-300
echo "--- custom barrier timing"; TSB_PCG_TIMING=1 timeout 300 python tools/prof.py --workload c2 --lin 1 --pcg 4000 --gn 2 2>&1 | grep "resident" | tail -1
