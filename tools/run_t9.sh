cd /root/repo
python -m pytest tests -m gpu -q -s 2>&1 | grep -E "trained|passed|failed|FAILED" | head -20 > gpurun_out/t9_tests.log
timeout 300 python bench.py --steps 5 --warmup 5 --no-cpu-baseline --in-flight 3 > gpurun_out/t9_bench_d3.json 2> gpurun_out/t9_bench_d3.err
timeout 300 python bench.py --steps 5 --warmup 5 --no-cpu-baseline > gpurun_out/t9_bench.json 2> gpurun_out/t9_bench.err
timeout 400 python bench.py --workload train256 --steps 3 --warmup 3 > gpurun_out/t9_train.json 2> gpurun_out/t9_train.err
cat gpurun_out/t9_tests.log; for f in t9_bench_d3 t9_bench t9_train; do cut -c1-200 gpurun_out/$f.json; done
