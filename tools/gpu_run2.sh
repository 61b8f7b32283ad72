set -x
timeout 1500 ncu --set full --clock-control none --import-source on -c 14 -o gpurun_out/r1_c4_full python tools/prof.py --workload c4 --lin 2 --pcg 5 > gpurun_out/ncu_c4_full.log 2>&1; echo "ncu rc=$?"
TSB_PCG_TIMING=1 timeout 600 python tools/prof.py --workload c4 --lin 2 --pcg 30 > gpurun_out/prof_c4_timing.log 2>&1
TSB_PCG_TIMING=1 timeout 600 python tools/prof.py --workload c2 --lin 2 --pcg 200 > gpurun_out/prof_c2_timing.log 2>&1
tail -3 gpurun_out/ncu_c4_full.log; cat gpurun_out/prof_c4_timing.log gpurun_out/prof_c2_timing.log
ls -la gpurun_out
