set -x
mkdir -p gpurun_out
timeout 400 python bench.py > gpurun_out/r2g_bench.json 2> gpurun_out/r2g_bench.err
timeout 300 python bench.py --impl reference --steps 5 --warmup 3 > gpurun_out/r2g_bench_reference_arm.json 2> gpurun_out/r2g_ref.err
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2g_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r2g_ncu_list.log 2>&1
timeout 400 ncu --set full --clock-control none --import-source on -k regex:k_tick_fused2 -s 6 -c 1 -o gpurun_out/r2g_k_tick_fused2 -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-otg > gpurun_out/r2g_ncu_full.log 2>&1
tail -2 gpurun_out/r2g_bench.json | cut -c1-600
