# quick GPU check of the update kernels: the GPU suite, then the C2 bench line (kernel times)
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -15
timeout 300 python bench.py --steps 20 --warmup 5 --no-port --no-c1 --no-dp-check > gpurun_out/bench_q.json 2> gpurun_out/bench_q.err
python - <<EOP
import json
for l in open("gpurun_out/bench_q.json"):
    if l.startswith("{"):
        d=json.loads(l); print(d["value"], d["ms_per_step"], d["e2e"]["value"], d.get("clocks"))
        for k in d["roofline"]["kernels"][:4]: print(k["key"], k["ms_per_launch"])
EOP
