timeout 300 python -m pytest tests/test_gpu_policy.py tests/test_gpu_large_batch.py -x -q -m gpu 2>&1 | tail -2
timeout 300 python bench.py --steps 20 --warmup 5 --no-port --no-c1 --no-dp-check > gpurun_out/bench_f.json 2> gpurun_out/bench_f.err
python - <<EOP
import json
for l in open("gpurun_out/bench_f.json"):
    if l.startswith("{"):
        d=json.loads(l); print(d["value"], d["ms_per_step"], d["e2e"]["value"])
        for k in d["roofline"]["kernels"][:2]: print(k["key"], k["ms_per_launch"])
EOP
timeout 600 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"mlp_tc_kernel<.int.2" -c 1 -f -o gpurun_out/prof_actor_f python tools/ncu_target.py > gpurun_out/ncu_actor_f.log 2>&1; tail -1 gpurun_out/ncu_actor_f.log
