python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r1c_gputests.log
python bench.py --steps 10 --warmup 3 > gpurun_out/r1c_bench.json 2> gpurun_out/r1c_bench.err
python tools/task_probe.py > gpurun_out/r1c_tasks.log 2>&1
N=16384 python tools/gpu_probe.py 2>&1 | tail -6 > gpurun_out/r1c_probe.log
cat gpurun_out/r1c_gputests.log gpurun_out/r1c_bench.json gpurun_out/r1c_tasks.log gpurun_out/r1c_probe.log
