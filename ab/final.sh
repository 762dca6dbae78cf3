cd /root/repo
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 300 python bench.py > gpurun_out/bench_final.log 2>&1; grep -h '{"metric"' gpurun_out/bench_final.log | tail -1 > gpurun_out/bench_final.json; cut -c1-200 gpurun_out/bench_final.json
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_final.log 2>&1; grep -h '"impl"' gpurun_out/bench_ref_final.log | tail -1 > gpurun_out/bench_ref_final.json; cut -c1-300 gpurun_out/bench_ref_final.json
K='regex:knn|filter|ransac|pack|gather|norms|bias|work|radius|scan|fix'
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 300 --csv --log-file gpurun_out/launches_final.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1; echo ncu1 $?
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 40 --csv --log-file gpurun_out/launches_radius.csv python tests/bench_radius.py 35 20000 8192 > gpurun_out/ncu_radius.log 2>&1; echo ncu2 $?
timeout 500 ncu --set full --clock-control none --import-source on -k regex:knn2_wide_kernel -c 1 -f -o gpurun_out/knn2_wide_final python tests/profile_knn.py 36 > gpurun_out/ncu_knn.log 2>&1; echo ncu3 $?
timeout 300 ncu --set full --clock-control none --import-source on -k regex:radius_kernel -c 1 -f -o gpurun_out/radius_final python tests/bench_radius.py 35 20000 8192 > gpurun_out/ncu_rad.log 2>&1; echo ncu4 $?
ls -la gpurun_out/*.ncu-rep
