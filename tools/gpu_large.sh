# GN timing of the large configs with per-phase cycles of the streamed PCG kernel
echo "--- c4 1 GN"; TSB_PCG_TIMING=1 timeout 900 python tools/prof.py --workload c4 --lin 1 --pcg 20000 --gn 1 2>&1 | grep "optimize\|streamed PCG" | cut -c1-400 | tail -2
echo "--- c5 2 GN"; TSB_PCG_TIMING=1 timeout 900 python tools/prof.py --workload c5 --lin 1 --pcg 20000 --gn 2 2>&1 | grep "optimize\|streamed PCG" | cut -c1-400 | tail -2
echo "--- c3 2 GN"; timeout 900 python tools/prof.py --workload c3 --lin 1 --pcg 20000 --gn 2 2>&1 | grep "optimize" | cut -c1-400 | tail -1
