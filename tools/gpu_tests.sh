# GPU test run: the allocation-free entry point alone in a fresh process, then the whole -m gpu suite (which ends with it)
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"; tail -4 gpurun_out/pytest_noalloc_alone.log | cut -c1-300
timeout 3000 python -m pytest tests -m gpu -x -q --durations=8 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest gpu rc=$?"; tail -40 gpurun_out/pytest_gpu.log | cut -c1-400
