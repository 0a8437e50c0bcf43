cd /root/repo
N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
timeout 600 $TR bench.py --gpus $N --workload stream11 --frames 64 > gpurun_out/n${N}_stream11.json 2> gpurun_out/n${N}_stream11.err
timeout 300 $TR bench.py --gpus $N --steps 10 --warmup 5 > gpurun_out/n${N}_frame10.json 2> gpurun_out/n${N}_frame10.err
if [ "$N" = "8" ]; then timeout 600 $TR bench.py --gpus $N --workload train256 --steps 5 --warmup 5 > gpurun_out/n${N}_train256.json 2> gpurun_out/n${N}_train256.err; fi
for f in stream11 frame10 train256; do cut -c1-200 gpurun_out/n${N}_$f.json 2>/dev/null; done
