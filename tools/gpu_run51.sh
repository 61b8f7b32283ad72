for v in 6 7 8 9; do echo "--- variant $v"; TSB_PS_VARIANT=$v TSB_PS_ONLY=a TSB_PCG_PHASED=1 timeout 600 python tools/prof.py --workload c3 --lin 1 --pcg 4 2>&1 | tail -1 | cut -c1-160; done
