# round 2 evidence: bench line (N=1), ncu launch list of the same command, ncu --set full of the top kernels
TAG=${1:-r2a}
timeout 1500 python bench.py --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_c3_n1.json 2> gpurun_out/${TAG}_bench_c3_n1.err; echo "bench rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/${TAG}_launches_bench_c3.csv python bench.py --steps 1 --warmup 1 --no-c4 --no-c2 --lin-reps 2 > gpurun_out/${TAG}_ncu_bench.log 2>&1; echo "launch list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_pcg_stream -c 1 -o gpurun_out/${TAG}_pcg_stream_c3 python tools/prof.py --workload c3 --lin 1 --pcg 8 > gpurun_out/${TAG}_ncu_ps3.log 2>&1; echo "ncu stream c3 rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_pcg_stream -c 1 -o gpurun_out/${TAG}_pcg_stream_c4 python tools/prof.py --workload c4 --lin 1 --pcg 6 > gpurun_out/${TAG}_ncu_ps4.log 2>&1; echo "ncu stream c4 rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k 'regex:k_linearize_stream|k_odom|k_lm_sum' -s 3 -c 3 -o gpurun_out/${TAG}_lin_c4 python tools/prof.py --workload c4 --lin 2 --pcg 0 > gpurun_out/${TAG}_ncu_lin4.log 2>&1; echo "ncu lin c4 rc=$?"
ls -la gpurun_out/*.ncu-rep
