echo "--- c3 phased streamed"; TSB_PCG_PHASED=1 timeout 600 python tools/prof.py --workload c3 --lin 1 --pcg 32 2>&1 | tail -1 | cut -c1-200
echo "--- c5 phased streamed"; TSB_PCG_PHASED=1 timeout 600 python tools/prof.py --workload c5 --lin 1 --pcg 32 2>&1 | tail -1 | cut -c1-200
echo "--- c4 phased streamed, no coarse"; TSB_COARSE=0 TSB_PCG_PHASED=1 timeout 600 python tools/prof.py --workload c4 --lin 1 --pcg 32 2>&1 | tail -1 | cut -c1-200
echo "--- c4 0.5 phased streamed"; TSB_PCG_PHASED=1 timeout 600 python tools/prof.py --workload c4 --scale 0.5 --lin 1 --pcg 32 2>&1 | tail -1 | cut -c1-200
