for B in 65536 1048576; do for M in 0 1 2; do
OSC_OTG_TMA=$M python bench.py --steps 20 --warmup 5 --no-cpu-baseline --batch $B 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); o=d['otg_on']; print($B,$M,round(d['ms_per_step'],4),round(o['ms_per_step'],4),round(o['k_otg_tick_x2_ms'],4))"
done; done
