run() { echo "smem=$1 grid=$2"; OSC_OTG_SLOW_SMEM=$1 OSC_OTG_SLOW_GRID=$2 timeout 200 python bench.py --steps 10 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); o=d['otg_on']; print('   recalc_ms', round(o['tick_all_robots_recalculate_ms'],4), 'otg_on_ms', round(o['ms_per_step'],4))"; }
run 0 0
run 120000 0
run 70000 0
run 0 148
run 0 296
run 120000 74
