cd /root/repo
V=${PCG_PROFILE_TAG:-r02}
# launch list (time-only pass) of one single-stream bench step; the bench must have exited 0 without ncu first
python bench.py --steps 1 --warmup 3 --no-cpu-baseline --in-flight 1 > gpurun_out/bench_pre_ncu.json 2> gpurun_out/bench_pre_ncu.err || exit 1
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file gpurun_out/launches_$V.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --in-flight 1 > gpurun_out/ncu_l_$V.log 2>&1
ls -la gpurun_out/launches_$V.csv
# --set full of the convolution / generative kernels (only when asked: the .ncu-rep stays on the box, the raw CSV travels)
if [ "$1" = "full" ]; then
  timeout 600 ncu --set full --clock-control none --import-source on -k "regex:k_conv_tc|k_gemm_first|k_conv_mma|k_cells_reduce|k_cells_kernel_map3|k_gather_sum" -c 29 -o /tmp/r02_${V}_conv -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --in-flight 1 > gpurun_out/ncu_full_$V.log 2>&1
  ncu -i /tmp/r02_${V}_conv.ncu-rep --page raw --csv > gpurun_out/r02_${V}_conv.raw.csv 2>/dev/null
  ls -la /tmp/r02_${V}_conv.ncu-rep gpurun_out/r02_${V}_conv.raw.csv
fi
