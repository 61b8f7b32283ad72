TSB_PCG_TIMING=1 timeout 600 python tools/prof.py --workload c4 --lin 1 --pcg 30 2>&1 | grep "streamed\|solve" | tail -3
