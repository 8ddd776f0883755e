#!/bin/bash
# quick GPU check: parity tests + the C2 bench line (kernel split); output under gpurun_out/<tag>.json
tag=${1:-quick}
python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/$tag.json 2> gpurun_out/$tag.err; tail -c 600 gpurun_out/$tag.err
python -c "
import json; d=json.load(open('gpurun_out/$tag.json')); print(d['value'], d['ms_per_step'], d['e2e']['value']); print({k:round(v['ms_per_step'],3) for k,v in d['kernel_ms_per_step'].items()})"
