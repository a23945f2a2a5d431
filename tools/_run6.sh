set -x
python bench.py --steps 2 --warmup 3 > gpurun_out/plain_bench.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r1h_launches.csv python bench.py --steps 2 --warmup 3 > gpurun_out/ncu_bench.log 2>&1
N=4736 LEVEL=1 WARPS=16 STEPS=3 python tools/ncu_one.py > gpurun_out/plain_one.log 2>&1 && N=4736 LEVEL=1 WARPS=16 STEPS=3 ncu --set full --clock-control none --import-source on -k regex:fly_step -s 3 -c 1 -o gpurun_out/r1h_prof python tools/ncu_one.py > gpurun_out/ncu_one.log 2>&1
tail -3 gpurun_out/plain_one.log gpurun_out/ncu_one.log; ls -la gpurun_out/
