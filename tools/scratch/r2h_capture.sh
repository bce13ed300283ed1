set -x
mkdir -p gpurun_out
for cfg in 4 3; do
  timeout 300 ncu --set full --clock-control none -k regex:k_tick_tail -s 8 -c 1 -o /tmp/r2h_tail_cfg$cfg -f python tools/bench_configs.py --only $cfg > gpurun_out/r2h_ncu_cfg$cfg.log 2>&1
  python tools/ncu_summary.py /tmp/r2h_tail_cfg$cfg.ncu-rep > gpurun_out/r2h_k_tick_tail_cfg$cfg.summary.txt 2>&1
  ncu -i /tmp/r2h_tail_cfg$cfg.ncu-rep --page raw --csv > gpurun_out/r2h_ncu_full_k_tick_tail_cfg$cfg.raw.csv 2>/dev/null
done
timeout 400 python tools/bench_configs.py > gpurun_out/r2h_configs.jsonl 2> gpurun_out/r2h_configs.err
timeout 300 python -m pytest tests/test_facade_cpp.py -m gpu -x -q 2>&1 | tail -5
tail -3 gpurun_out/r2h_configs.jsonl | cut -c1-300
du -sh gpurun_out
