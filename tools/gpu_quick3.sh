# GPU suite, then C4 / C5 step times
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
for c in c4 c5; do timeout 600 python bench.py --config $c --steps 5 --warmup 3 --no-cpu --e2e-steps 0 --no-c1 --no-dp-check 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('$c', d['value'], d['ms_per_step'])"; done
