timeout 200 python bench.py --steps 10 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); o=d['otg_on']; print('   recalc_ms', round(o['tick_all_robots_recalculate_ms'],4), 'otg_on_ms', round(o['ms_per_step'],4))"
timeout 400 python -m pytest tests/test_otg_gpu.py -m gpu -x -q 2>&1 | tail -2
