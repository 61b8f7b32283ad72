echo "--- always refresh"; TSB_COARSE_MAX_AGE=1 timeout 300 python tools/prof.py --workload c2 --lin 1 --pcg 4000 --gn 50 2>&1 | tail -1 | cut -c1-900
echo "--- max age 3"; TSB_COARSE_MAX_AGE=3 timeout 300 python tools/prof.py --workload c2 --lin 1 --pcg 4000 --gn 50 2>&1 | tail -1 | cut -c1-200
echo "--- max age 2"; TSB_COARSE_MAX_AGE=2 timeout 300 python tools/prof.py --workload c2 --lin 1 --pcg 4000 --gn 50 2>&1 | tail -1 | cut -c1-200
echo "--- default"; timeout 300 python tools/prof.py --workload c2 --lin 1 --pcg 4000 --gn 50 2>&1 | tail -1 | cut -c1-900
