cd /root/repo
run() { echo "=== $*"; env "$@" python tools/bench_conv.py --heads 2>&1 | grep -v "^$" | tail -7; }
run A=0
run PCG_TC_CA=1
run PCG_TC_CA=1 PCG_TC_CARVEOUT=50
run PCG_TC_CA=1 PCG_TC_CARVEOUT=25
run PCG_TC_CA=1 PCG_TC_CARVEOUT=50 PCG_TC_NARROW_KB=30
run PCG_TC_CA=3
run PCG_TC_CA=3 PCG_TC_CARVEOUT=50
run PCG_TC_CA=3 PCG_TC_CARVEOUT=50 PCG_NARROW=tc
run PCG_NARROW=tc
