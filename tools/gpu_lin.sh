# linearize time at C4 / C3 + per-phase PCG cycles + the bit-identity tests of the streamed kernels
for w in c4 c3; do echo "--- $w"; TSB_PCG_TIMING=1 timeout 900 python tools/prof.py --workload $w --lin 6 --pcg 20000 --gn 1 2>&1 | grep "linearize\|streamed PCG\|optimize" | cut -c1-400 | tail -6; done
timeout 900 python -m pytest tests -m gpu -x -q -k "bit_identical or streamed or reproducible or multilevel or full_size_solve" 2>&1 | tail -3
