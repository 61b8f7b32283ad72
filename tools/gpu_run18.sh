for a in 0 1 2 3 4 7 15; do echo "--- ABL=$a"; TSB_LS_ABL=$a timeout 600 python tools/prof.py --workload c4 --lin 4 --pcg 0 2>&1 | tail -1; done
