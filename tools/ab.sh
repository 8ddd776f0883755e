# A/B of two builds of the library on ONE box: tools/ab.sh path/libA.so path/libB.so  (bench kernel times of each, twice)
for rep in 1 2; do
  for v in "$1" "$2"; do
    cp "$v" mappo_b200/libmappo_b200.so
    timeout 300 python bench.py --steps 20 --warmup 5 --no-port --no-c1 --no-dp-check 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('$v', d['ms_per_step'], [round(k['ms_per_launch'],4) for k in d['roofline']['kernels'][:4]])"
  done
done
