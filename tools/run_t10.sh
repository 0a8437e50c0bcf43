cd /root/repo
python -m pytest tests -m gpu -q 2>&1 | tail -15 > gpurun_out/t10_tests.log
timeout 300 python bench.py --workload stream11 --frames 8 > gpurun_out/t10_stream8_batched.json 2> gpurun_out/t10_stream8_batched.err
timeout 300 python bench.py --workload stream11 --frames 8 --serial-blocks > gpurun_out/t10_stream8_serial.json 2> gpurun_out/t10_stream8_serial.err
timeout 300 python bench.py --steps 5 --warmup 5 --no-cpu-baseline > gpurun_out/t10_bench.json 2> gpurun_out/t10_bench.err
tail -3 gpurun_out/t10_tests.log; for f in t10_stream8_batched t10_stream8_serial t10_bench; do cut -c1-160 gpurun_out/$f.json; tail -1 gpurun_out/$f.err | cut -c1-200; done
