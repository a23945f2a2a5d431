for t in walk_imitation walk_on_ball flight_imitation; do
TASK=$t python tools/variant_probe.py >> gpurun_out/r1t_variants.log 2>&1
done
cat gpurun_out/r1t_variants.log
python -m pytest tests -m gpu -q 2>&1 | tail -3
