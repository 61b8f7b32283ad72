# per-phase cycles of the streamed PCG at C3 / C4, a 50-iteration C3 request (PCG iterations / ms per GN iteration), parity subset
for w in c3 c4; do echo "--- $w"; TSB_PCG_TIMING=1 timeout 900 python tools/prof.py --workload $w --lin 2 --pcg 20000 --gn 2 2>&1 | grep -v "^$" | grep -v "k_ml_top" | cut -c1-500 | tail -5; done
echo "--- c3 x 50"; timeout 600 python tools/prof.py --workload c3 --lin 0 --pcg 0 --gn 50 2>&1 | tail -1 | cut -c1-1500
timeout 900 python -m pytest tests -m gpu -x -q -k "multilevel or full_size_solve or streamed or reproducible or golden or optimize_matches or more_edges or load_soa" 2>&1 | tail -3
