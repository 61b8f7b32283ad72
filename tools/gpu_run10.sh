timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -7 gpurun_out/pytest_gpu.log
timeout 600 python tools/prof.py --workload c4 --lin 6 --pcg 0 2>&1 | tail -3
TSB_PCG_TIMING=1 timeout 600 python tools/prof.py --workload c2 --lin 1 --pcg 4000 --gn 3 2>&1 | tail -12 | cut -c1-300
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/launches_c4_lin.csv python tools/prof.py --workload c4 --lin 3 --pcg 0 > /dev/null 2>&1; grep -v "^==" gpurun_out/launches_c4_lin.csv | awk -F'","' 'NR>1{print $5, $NF}' | tail -12
