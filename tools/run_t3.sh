cd /root/repo
python -m pytest tests/test_gpu_conv.py -m gpu -q -x -k "weight_gradient_on_tensor_cores" 2>&1 | tail -70 > gpurun_out/t3_dw.log
python -m pytest tests -m gpu -q 2>&1 | tail -30 > gpurun_out/t3_tests.log
timeout 600 python tools/train_profile.py 2 > gpurun_out/t3_trainprof.log 2>&1
tail -5 gpurun_out/t3_dw.log; tail -3 gpurun_out/t3_tests.log
