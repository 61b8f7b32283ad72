# quick single-GPU check after a kernel change: per-phase cycles of the streamed PCG at C3 / C4 + a few parity tests
for w in c3 c4; do echo "--- $w"; TSB_PCG_TIMING=1 timeout 900 python tools/prof.py --workload $w --lin 2 --pcg 20000 --gn 2 2>&1 | grep -v "^$" | cut -c1-500 | tail -6; done
timeout 900 python -m pytest tests -m gpu -x -q -k "multilevel or full_size_solve or streamed or reproducible or golden or optimize_matches" 2>&1 | tail -3
