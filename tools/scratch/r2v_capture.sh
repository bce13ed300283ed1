set -x
mkdir -p gpurun_out
timeout 200 ncu --set full --clock-control none -k regex:k_tick_fused2 -s 8 -c 1 -o /tmp/r2v_s7 -f python tools/bench_configs.py --only st7 > gpurun_out/r2v_ncu_s7.log 2>&1
python tools/ncu_summary.py /tmp/r2v_s7.ncu-rep > gpurun_out/r2v_k_tick_fused2_single7.summary.txt 2>&1
timeout 200 ncu --set full --clock-control none -k regex:k_tick_joint -s 8 -c 1 -o /tmp/r2v_jt -f python tools/bench_configs.py --only jt > gpurun_out/r2v_ncu_jt.log 2>&1
python tools/ncu_summary.py /tmp/r2v_jt.ncu-rep > gpurun_out/r2v_k_tick_joint_puma.summary.txt 2>&1
head -30 gpurun_out/r2v_k_tick_fused2_single7.summary.txt; head -30 gpurun_out/r2v_k_tick_joint_puma.summary.txt
