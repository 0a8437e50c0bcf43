cd /root/repo
# source-level (SASS + CUDA-C line) profile of ONE launch of a named kernel inside a bench step
#   usage: tools/profile_source_kernel.sh <kernel regex> <launch-skip> <tag>
K=${1:-k_cells_reduce}; SKIP=${2:-1}; TAG=${3:-reduce}
python bench.py --steps 1 --warmup 3 --no-cpu-baseline --in-flight 1 > gpurun_out/src_pre.json 2> gpurun_out/src_pre.err || exit 1
timeout 300 ncu --set full --section SourceCounters --clock-control none --import-source on -k "regex:$K" --launch-skip $SKIP -c 1 -o /tmp/r02_src_$TAG -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --in-flight 1 > gpurun_out/ncu_src_$TAG.log 2>&1
ncu -i /tmp/r02_src_$TAG.ncu-rep --page source --csv --print-source sass,cuda > gpurun_out/r02_src_${TAG}_both.csv 2>/dev/null
ncu -i /tmp/r02_src_$TAG.ncu-rep --page raw --csv > gpurun_out/r02_src_${TAG}_raw.csv 2>/dev/null
ls -la gpurun_out/r02_src_${TAG}_*
