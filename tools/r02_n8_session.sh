# round 2, 8-GPU session: multi-GPU tests, torchrun bench at N = 8 and 4, config 4 (row-sharded) both ways
set -x
mkdir -p gpurun_out
nproc; nvidia-smi -L | wc -l
timeout 900 python -m pytest tests/test_gpu_multi.py -q -m gpu --timeout 600 2>&1 | tail -4
for N in 8 4; do
NCCL_DEBUG=INFO timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N bench.py --gpus $N --steps 100 --warmup 3 > gpurun_out/r02_bench_n$N.json 2> gpurun_out/r02_bench_n$N.err
grep -c "NCCL INFO" gpurun_out/r02_bench_n$N.err; grep "nranks" gpurun_out/r02_bench_n$N.err | head -2 | cut -c1-200
cut -c1-160 gpurun_out/r02_bench_n$N.json
done
timeout 900 python tools/bench_row_sharded.py --log2-rows 27 --candidates 160000 --steps 10 2>/dev/null | tail -1 > gpurun_out/r02_config4_2p27_fp32.json; cat gpurun_out/r02_config4_2p27_fp32.json
timeout 900 python tools/bench_row_sharded.py --log2-rows 30 --candidates 160000 --steps 10 --table-dtype f16 --skip-replicated 2>/dev/null | tail -1 > gpurun_out/r02_config4_2p30_f16.json; cat gpurun_out/r02_config4_2p30_f16.json
