# spread of the PCG phases over the blocks (TSB_PCG_TIMING_BLOCKS=1) + linearize time
for w in c3 c4; do echo "--- $w"; TSB_PCG_TIMING=1 TSB_PCG_TIMING_BLOCKS=1 timeout 900 python tools/prof.py --workload $w --lin 4 --pcg 20000 --gn 1 2>&1 | grep "linearize\|streamed PCG\|over .* blocks\|optimize" | cut -c1-330 | tail -7; done
