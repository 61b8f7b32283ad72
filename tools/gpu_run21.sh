for v in A B C D; do cp tmp_libs/lib$v.so toyslam_b200/libtoyslam_b200.so; echo "--- variant $v c4"; timeout 600 python tools/prof.py --workload c4 --lin 5 --pcg 0 2>&1 | tail -2; done
