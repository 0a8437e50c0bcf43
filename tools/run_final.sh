cd /root/repo
python -m pytest tests -m gpu -q 2>&1 | tail -6 > gpurun_out/f_tests.log
timeout 300 python bench.py > gpurun_out/f_bench.json 2> gpurun_out/f_bench.err
timeout 300 python bench.py --steps 10 --warmup 5 --no-cpu-baseline --topk-from encoder > gpurun_out/f_bench_enc.json 2> gpurun_out/f_bench_enc.err
timeout 300 python bench.py --workload block128 --steps 10 --warmup 5 > gpurun_out/f_block128.json 2> gpurun_out/f_block128.err
timeout 300 python bench.py --workload train256 --steps 5 --warmup 5 > gpurun_out/f_train.json 2> gpurun_out/f_train.err
PCG_PROFILE_TAG=v17 bash tools/profile_step.sh full > gpurun_out/f_profile.log 2>&1
tail -3 gpurun_out/f_tests.log; for f in f_bench f_bench_enc f_block128 f_train; do cut -c1-150 gpurun_out/$f.json; done; tail -2 gpurun_out/f_profile.log
