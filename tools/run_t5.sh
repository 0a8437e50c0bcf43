cd /root/repo
python -m pytest tests/test_gpu_model.py -m gpu -q -s -k trained 2>&1 | grep -E "trained|passed|failed|assert|Error" | head -20 > gpurun_out/t5_trained.log
python tools/host_profile.py block128 > gpurun_out/t5_host_block128.log 2>&1
python tools/host_profile.py frame10 > gpurun_out/t5_host_frame10.log 2>&1
cat gpurun_out/t5_trained.log
