cd /root/repo
python -m pytest tests -m gpu -q 2>&1 | tail -60 > gpurun_out/t2_tests.log
timeout 300 python bench.py --steps 5 --warmup 5 > gpurun_out/t2_bench.json 2> gpurun_out/t2_bench.err
timeout 200 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --topk-from encoder > gpurun_out/t2_bench_enc.json 2> gpurun_out/t2_bench_enc.err
timeout 200 python bench.py --workload block128 --steps 5 --warmup 5 > gpurun_out/t2_block128.json 2> gpurun_out/t2_block128.err
timeout 400 python bench.py --workload train256 --steps 3 --warmup 3 > gpurun_out/t2_train.json 2> gpurun_out/t2_train.err
timeout 400 python bench.py --workload stream11 --frames 8 > gpurun_out/t2_stream8.json 2> gpurun_out/t2_stream8.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/t2_ref.json 2> gpurun_out/t2_ref.err
tail -3 gpurun_out/t2_tests.log; for f in t2_bench t2_bench_enc t2_block128 t2_train t2_stream8 t2_ref; do echo $f; cut -c1-300 gpurun_out/$f.json; tail -2 gpurun_out/$f.err; done
