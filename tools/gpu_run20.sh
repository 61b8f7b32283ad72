timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_gpu.log
echo "--- c4 v2"; timeout 600 python tools/prof.py --workload c4 --lin 6 --pcg 0 2>&1 | tail -2 | tee gpurun_out/prof_c4_planes.log
echo "--- c4 ws"; TSB_STREAM_WS=1 timeout 600 python tools/prof.py --workload c4 --lin 6 --pcg 0 2>&1 | tail -2
echo "--- c4 unit kernel"; TSB_NO_STREAM=1 timeout 600 python tools/prof.py --workload c4 --lin 4 --pcg 0 2>&1 | tail -1
echo "--- c3 v2"; timeout 600 python tools/prof.py --workload c3 --lin 6 --pcg 0 2>&1 | tail -2
echo "--- c4 pcg"; timeout 900 python tools/prof.py --workload c4 --lin 1 --pcg 40 2>&1 | tail -3 | cut -c1-400
echo "--- bench"; timeout 900 python bench.py > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "bench rc=$?"; cut -c1-700 gpurun_out/bench_c2.json
