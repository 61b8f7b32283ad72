timeout 900 python -m pytest tests -m gpu -x -q -k "more_edges_than_a_chunk or degenerate" > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/pytest_gpu.log | cut -c1-220
