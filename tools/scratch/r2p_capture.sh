set -x
mkdir -p gpurun_out
timeout 300 ncu --set full --clock-control none -k regex:k_tick_fused3 -s 8 -c 1 -o /tmp/r2p_f3 -f python tools/bench_configs.py --only 3 > gpurun_out/r2p_ncu_f3.log 2>&1
python tools/ncu_summary.py /tmp/r2p_f3.ncu-rep > gpurun_out/r2p_k_tick_fused3_cfg3.summary.txt 2>&1
timeout 300 ncu --set full --clock-control none -k regex:k_tick_fused2 -s 8 -c 1 -o /tmp/r2p_f2p -f python tools/bench_configs.py --only st > gpurun_out/r2p_ncu_f2p.log 2>&1
python tools/ncu_summary.py /tmp/r2p_f2p.ncu-rep > gpurun_out/r2p_k_tick_fused2_puma.summary.txt 2>&1
cat gpurun_out/r2p_k_tick_fused3_cfg3.summary.txt | head -24
