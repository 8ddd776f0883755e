# end-of-round check on one GPU: GPU suite, smoke, both bench arms, supplementary configs
set -x
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -2
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err; echo bench rc=$?
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/final_ref.json 2> gpurun_out/final_ref.err; echo ref rc=$?
for c in c3 c3r c4 c5; do timeout 600 python bench.py --config $c --steps 5 --warmup 3 --no-cpu --no-c1 --no-dp-check > gpurun_out/final_$c.json 2> gpurun_out/final_$c.err; echo $c rc=$?; done
timeout 300 python -m pytest tests/test_zz_no_alloc.py -q -m gpu 2>&1 | tail -1
