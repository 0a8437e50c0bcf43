cd /root/repo
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29517"
timeout 300 $TR bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/n8c_frame10_pin.json 2> gpurun_out/n8c_frame10_pin.err
PCG_PIN=0 timeout 300 $TR bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/n8c_frame10_nopin.json 2> gpurun_out/n8c_frame10_nopin.err
lscpu | grep -E "Model name|Thread|Core|Socket|^CPU\(s\)" > gpurun_out/n8c_lscpu.txt
cut -c1-160 gpurun_out/n8c_frame10_pin.json gpurun_out/n8c_frame10_nopin.json; cat gpurun_out/n8c_lscpu.txt
