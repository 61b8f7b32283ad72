timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -7 gpurun_out/pytest_gpu.log
timeout 600 python tools/prof.py --workload c4 --lin 6 --pcg 0 2>&1 | tee gpurun_out/prof_c4_stream.log
timeout 600 python tools/prof.py --workload c3 --lin 6 --pcg 0 2>&1 | tail -3 | tee gpurun_out/prof_c3_stream.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_linearize_stream -s 1 -c 1 -o gpurun_out/r1_lin_stream2_c4 python tools/prof.py --workload c4 --lin 3 --pcg 0 > gpurun_out/ncu_lin_stream.log 2>&1; echo "ncu rc=$?"
