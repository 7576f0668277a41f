#!/bin/bash
# push_many: the tape cut (tree) against the bundled segment cut (MOF_SEQ_CUT=bundles) and the uniform one (MOF_SEQ_UNIFORM=1)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_round2.py tests/test_gpu_generic_sizes.py tests/test_gpu_steady_state.py -x -q -m gpu -k "push_many or stream" 2>&1 | tail -3
MOF_SEQ_CUT=bundles timeout 900 python -m pytest tests/test_gpu_round2.py -x -q -m gpu -k "push_many" 2>&1 | tail -2
{
for s in 8 16 32 64; do
  for rep in 1 2; do
  echo "tape, $s streams x 64 frames: $(timeout 300 python bench.py --workload streams --streams $s --steps 5 --warmup 3 --no-e2e 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'])")"
  echo "bundles, $s streams x 64 frames: $(MOF_SEQ_CUT=bundles timeout 300 python bench.py --workload streams --streams $s --steps 5 --warmup 3 --no-e2e 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'])")"
  done
done
} > gpurun_out/r5h_ab_seq_tape.txt 2>&1
cat gpurun_out/r5h_ab_seq_tape.txt
