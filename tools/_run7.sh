python tools/drift_report.py > gpurun_out/r1_drift.jsonl 2> gpurun_out/r1_drift.err; tail -3 gpurun_out/r1_drift.err; cat gpurun_out/r1_drift.jsonl
python bench.py > gpurun_out/r1_bench_final.json 2> gpurun_out/r1_bench_final.err; tail -2 gpurun_out/r1_bench_final.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r1_bench_reference_arm.json 2> gpurun_out/r1_ref.err; tail -2 gpurun_out/r1_ref.err; cut -c1-300 gpurun_out/r1_bench_reference_arm.json
python bench.py --steps 2 --warmup 3 > gpurun_out/plain_bench4.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r1_launches_final.csv python bench.py --steps 2 --warmup 3 > gpurun_out/ncu_l.log 2>&1
python bench.py --steps 2 --warmup 3 > gpurun_out/plain_bench5.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:fly_step -s 4 -c 1 -o gpurun_out/r1_final_prof python bench.py --steps 2 --warmup 3 > gpurun_out/ncu_f.log 2>&1
ls -la gpurun_out/r1_final_prof.ncu-rep gpurun_out/r1_launches_final.csv
