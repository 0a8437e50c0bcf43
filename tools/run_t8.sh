cd /root/repo
for d in 1 2 3 4; do timeout 300 python bench.py --workload stream11 --frames 8 --in-flight $d > gpurun_out/t8_stream8_d$d.json 2> gpurun_out/t8_stream8_d$d.err; done
python -m pytest tests/test_gpu_conv.py tests/test_gpu_model.py -m gpu -q 2>&1 | tail -5 > gpurun_out/t8_tests.log
timeout 300 python bench.py --steps 5 --warmup 5 --no-cpu-baseline > gpurun_out/t8_bench.json 2> gpurun_out/t8_bench.err
timeout 300 python tools/time_dropin.py frame10 > gpurun_out/r02_dropin_timing.json 2> gpurun_out/t8_dropin.err
for d in 1 2 3 4; do cut -c1-120 gpurun_out/t8_stream8_d$d.json; done; tail -2 gpurun_out/t8_tests.log; cat gpurun_out/r02_dropin_timing.json
