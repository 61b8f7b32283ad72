timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_sb_build -c 1 -o gpurun_out/r1_sb_build_c2_v2 python tools/prof.py --workload c2 --lin 1 --pcg 20000 --gn 0 > gpurun_out/ncu_sbb.log 2>&1; echo "ncu rc=$?"; tail -1 gpurun_out/ncu_sbb.log
timeout 900 python -m pytest tests -m gpu -x -q -k "resident or full_size or reproducible" 2>&1 | tail -2
