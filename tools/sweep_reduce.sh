cd /root/repo
for v in 1 2; do echo "=== VPL=$v"; PCG_REDUCE_VPL=$v python tools/bench_cells.py 2>&1 | tail -4; done
