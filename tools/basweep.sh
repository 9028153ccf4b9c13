for g in 148 111 96 74 64 48 37 24; do echo "BA grid $g"; WVB200_BA_GRID=$g python bench.py --no-cpu --steps 20 --warmup 5 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['ms_breakdown'])"; done
