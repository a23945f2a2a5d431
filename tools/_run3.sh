for b in 0 1; do for t in walk_imitation walk_on_ball; do
FLYB_BINNING=$b TASK=$t WARM=12 python tools/variant_probe.py >> gpurun_out/r1r_variants.log 2>&1
done; done
cat gpurun_out/r1r_variants.log
