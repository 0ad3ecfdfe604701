# what the driver runs at round end, on one GPU
set -x
python -m pytest tests/ -x -q -m gpu 2>&1 | tail -3
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
(time python bench.py --impl reference --gpus 1 --steps 200 --warmup 3) > gpurun_out/r02_final_ref.json 2> gpurun_out/r02_final_ref.err; tail -3 gpurun_out/r02_final_ref.err
(time python bench.py --gpus 1 --steps 200 --warmup 3) > gpurun_out/r02_final_bench.json 2> gpurun_out/r02_final_bench.err; tail -3 gpurun_out/r02_final_bench.err
wc -l gpurun_out/r02_final_ref.json gpurun_out/r02_final_bench.json
