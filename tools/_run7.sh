python bench.py --steps 2 --warmup 3 > gpurun_out/plain_bench2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:fly_step -s 4 -c 1 -o gpurun_out/r1h_bench_prof python bench.py --steps 2 --warmup 3 > gpurun_out/ncu_bench2.log 2>&1
ls -la gpurun_out/*.ncu-rep
