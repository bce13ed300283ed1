set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -3 > gpurun_out/r2k_pytest_gpu.txt
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2k_smoke.txt 2>&1
timeout 400 python bench.py > gpurun_out/r2k_bench.json 2> gpurun_out/r2k_bench.err
timeout 300 python bench.py --impl reference --steps 5 --warmup 3 > gpurun_out/r2k_bench_reference_arm.json 2> gpurun_out/r2k_ref.err
timeout 400 python tools/bench_configs.py > gpurun_out/r2k_configs.jsonl 2> gpurun_out/r2k_configs.err
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2k_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r2k_ncu_list.log 2>&1
cat gpurun_out/r2k_pytest_gpu.txt gpurun_out/r2k_smoke.txt
