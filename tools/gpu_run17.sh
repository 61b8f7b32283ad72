timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_gpu.log
echo "--- c4 ws"; timeout 600 python tools/prof.py --workload c4 --lin 6 --pcg 0 2>&1 | tail -3 | tee gpurun_out/prof_c4_ws.log
echo "--- c4 v2"; TSB_STREAM_WS=0 timeout 600 python tools/prof.py --workload c4 --lin 6 --pcg 0 2>&1 | tail -3
echo "--- c3 ws"; timeout 600 python tools/prof.py --workload c3 --lin 6 --pcg 0 2>&1 | tail -3
echo "--- c3 v2"; TSB_STREAM_WS=0 timeout 600 python tools/prof.py --workload c3 --lin 6 --pcg 0 2>&1 | tail -3
echo "--- bench"; timeout 900 python bench.py > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "bench rc=$?"; cut -c1-1500 gpurun_out/bench_c2.json
