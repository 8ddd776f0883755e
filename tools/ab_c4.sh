# A/B of two builds on ONE box, configuration C4 (and the C2 line as a guard): tools/ab_c4.sh libA.so libB.so
for v in "$1" "$2" "$1" "$2"; do
  cp "$v" mappo_b200/libmappo_b200.so
  for c in c4; do timeout 600 python bench.py --config $c --steps 5 --warmup 3 --no-cpu --e2e-steps 0 --no-c1 --no-dp-check 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('$v $c', d['ms_per_step'])"; done
done
cp "$2" mappo_b200/libmappo_b200.so
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --e2e-steps 0 --no-c1 --no-dp-check 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('$2 c2', d['ms_per_step'])"
