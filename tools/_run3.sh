for v in base cg; do for t in walk_imitation; do TASK=$t FLYB_LIB_PATH=$PWD/flybody_b200/_exp/libflyb_$v.so python tools/variant_probe.py >> gpurun_out/r1i_variants.log 2>&1; done; done
cat gpurun_out/r1i_variants.log
