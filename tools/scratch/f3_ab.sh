for R in 0 168 255; do
echo "OSC_F3_REGS=$R"
OSC_F3_REGS=$R timeout 200 python tools/bench_configs.py --only 3 2>&1 | tail -1 | cut -c1-200
done
timeout 200 python -m pytest tests/test_fused3_gpu.py tests/test_config3_force_popc_gpu.py -m gpu -x -q 2>&1 | tail -2
timeout 120 ncu --nvtx --nvtx-include "osc_tick/" --metrics gpu__time_duration.sum --clock-control none -c 6 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-otg 2>&1 | grep -i "k_tick\|nvtx\|==PROF==" | head -12
