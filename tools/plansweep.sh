for g in 148 96 64 48 32 16; do echo "plan grid $g"; WVB200_PLAN_GRID=$g python bench.py --no-cpu --steps 20 --warmup 5 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['ms_breakdown'])"; done
