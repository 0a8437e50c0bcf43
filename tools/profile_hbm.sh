cd /root/repo
# ncu --set full of the HBM-bound integer / entropy kernels of one bench step (hash, kernel map, compaction, top-k, entropy model, GDN).
# Only the raw CSV travels back (the .ncu-rep of ~150 launches exceeds gpurun's 64 MiB return limit).
timeout 420 ncu --set full --clock-control none \
  -k "regex:k_kernel_map|k_hash_insert|k_insert_|k_gc_|k_eb_|k_quantize|k_dequantize|k_features_at|k_gather_rows|pcg_k_scatter|pcg_k_flag_count|k_topk_|k_remap_table|k_gdn|k_im2col|k_cells_parent|k_cells_emit|k_cells_mask" \
  -c 170 -o /tmp/r01_v13_hbm -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --in-flight 1 > gpurun_out/ncu_hbm13.log 2>&1
ncu -i /tmp/r01_v13_hbm.ncu-rep --page raw --csv > gpurun_out/r01_v13_hbm.raw.csv 2>/dev/null
ls -la /tmp/r01_v13_hbm.ncu-rep gpurun_out/r01_v13_hbm.raw.csv
