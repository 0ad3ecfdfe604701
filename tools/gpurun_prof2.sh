P="python bench.py --no-e2e --no-cpu-baseline --steps 3 --warmup 3 --requests-per-step 64"
$P > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"gather_pool|tower_chain" -c 60 --csv --log-file gpurun_out/r01_e_launches.csv $P > gpurun_out/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"gather_pool|tower_chain" -s 6 -c 2 -o gpurun_out/r01_e_prof $P > gpurun_out/ncu_f.log 2>&1
tail -n 3 gpurun_out/ncu_l.log gpurun_out/ncu_f.log
python bench.py --steps 100 --warmup 5 > gpurun_out/r01_e_bench.json 2> gpurun_out/r01_e_bench.err; tail -c 1800 gpurun_out/r01_e_bench.json
