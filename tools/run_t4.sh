cd /root/repo
python -m pytest tests -m gpu -q 2>&1 | tail -40 > gpurun_out/t4_tests.log
timeout 600 python tools/train_profile.py 8 > gpurun_out/t4_trainprof.log 2>&1
timeout 400 python bench.py --workload train256 --steps 3 --warmup 3 > gpurun_out/t4_train.json 2> gpurun_out/t4_train.err
timeout 300 python bench.py --steps 5 --warmup 5 --no-cpu-baseline > gpurun_out/t4_bench.json 2> gpurun_out/t4_bench.err
tail -3 gpurun_out/t4_tests.log; cut -c1-400 gpurun_out/t4_train.json; grep "step wall" gpurun_out/t4_trainprof.log
