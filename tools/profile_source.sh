cd /root/repo
# source-level (SASS + CUDA-C line) profile of ONE launch: the level-2 first head layer with the fused projection
# (k_conv_tc_mt<2,128,false,64>, the 16th k_conv_tc_mt launch of a bench step)
timeout 300 ncu --set full --section SourceCounters --clock-control none --import-source on -k "regex:k_conv_tc_mt" --launch-skip ${1:-15} -c 1 -o /tmp/r01_src -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --in-flight 1 > gpurun_out/ncu_src.log 2>&1
ncu -i /tmp/r01_src.ncu-rep --page source --csv > gpurun_out/r01_src_sass.csv 2>/dev/null; ncu -i /tmp/r01_src.ncu-rep --page source --csv --print-source sass,cuda > gpurun_out/r01_src_both.csv 2>/dev/null
ncu -i /tmp/r01_src.ncu-rep --page raw --csv > gpurun_out/r01_src_raw.csv 2>/dev/null
ls -la gpurun_out/r01_src_*
