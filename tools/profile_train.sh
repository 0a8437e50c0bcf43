cd /root/repo
# ncu --set full of the weight-gradient kernels of one training step (8 blocks of 256^3); the plain run must exit 0 first
python tools/train_profile.py 8 > gpurun_out/trainprof_pre.log 2>&1 || exit 1
timeout 500 ncu --set full --clock-control none --import-source on -k "regex:k_conv_bwd_weight" -c 24 -o /tmp/r02_train_bwd -f python tools/train_profile.py 8 > gpurun_out/ncu_train_bwd.log 2>&1
ncu -i /tmp/r02_train_bwd.ncu-rep --page raw --csv > gpurun_out/r02_train_bwd.raw.csv 2>/dev/null
ls -la gpurun_out/r02_train_bwd.raw.csv; head -24 gpurun_out/trainprof_pre.log
