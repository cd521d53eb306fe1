#!/bin/bash
# Round 2, GPU call H (1 GPU): categorical DMMA kernel with two tiles per iteration: parity, then A/B.
O=gpurun_out/r02h; mkdir -p $O
( time timeout 900 python -m pytest tests/test_categorical_mma_gpu.py tests/test_glm_gpu.py tests/test_stress_gpu.py -m gpu -x -q -k "categorical or widened" ) > $O/pytest_cat.log 2>&1; echo "pytest cat rc=$?" | tee -a $O/rc.txt
for pair in 2 0; do
  STAN_CUDA_CAT_PAIR=$pair timeout 900 python tools/ab.py --widened --cat-kernel mma --only categorical > $O/ab_cat_pair$pair.jsonl 2> $O/ab_cat_pair$pair.err; echo "ab pair=$pair rc=$?" | tee -a $O/rc.txt
done
tail -3 $O/pytest_cat.log; cat $O/rc.txt
paste -d'\n' $O/ab_cat_pair1.jsonl $O/ab_cat_pair0.jsonl | cut -c1-150
