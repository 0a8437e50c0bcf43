cd /root/repo
python bench.py --steps 1 --warmup 3 --no-cpu-baseline --in-flight 1 > gpurun_out/bench_pre_ncu.json 2> gpurun_out/bench_pre_ncu.err || exit 1
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file gpurun_out/launches_v12.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --in-flight 1 > gpurun_out/ncu_l12.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k "regex:k_conv_tc|k_gemm_first|k_conv_mma|k_cells_reduce|k_cells_kernel_map3" -c 29 -o gpurun_out/r01_v12_conv -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --in-flight 1 > gpurun_out/ncu_full12.log 2>&1
ncu -i gpurun_out/r01_v12_conv.ncu-rep --page raw --csv > gpurun_out/r01_v12_conv.raw.csv 2>/dev/null
ls -la gpurun_out/r01_v12_conv* gpurun_out/launches_v12.csv
