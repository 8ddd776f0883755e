# quick GPU check of the update kernels: parity subsets, then the C2 bench line (kernel times)
timeout 300 python -m pytest tests/test_gpu_policy.py tests/test_gpu_large_batch.py -x -q -m gpu 2>&1 | tail -2
timeout 300 python bench.py --steps 20 --warmup 5 --no-port --no-c1 --no-dp-check > gpurun_out/bench_q.json 2> gpurun_out/bench_q.err
python - <<EOP
import json
for l in open("gpurun_out/bench_q.json"):
    if l.startswith("{"):
        d=json.loads(l); print(d["value"], d["ms_per_step"], d["e2e"]["value"])
        for k in d["roofline"]["kernels"][:2]: print(k["key"], k["ms_per_launch"])
EOP
