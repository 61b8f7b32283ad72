timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_c2_sb.csv python tools/prof.py --workload c2 --lin 1 --pcg 20000 --gn 2 > gpurun_out/ncu_sb.log 2>&1; echo "ncu rc=$?"; grep k_sb_build gpurun_out/launches_c2_sb.csv | cut -d, -f5,12- | head -3
echo "--- c2 50 GN adaptive"; timeout 300 python tools/prof.py --workload c2 --lin 1 --pcg 20000 --gn 50 2>&1 | grep "optimize" | cut -c1-200
timeout 900 python -m pytest tests -m gpu -x -q -k "resident or full_size or reproducible or oracle" 2>&1 | tail -2
