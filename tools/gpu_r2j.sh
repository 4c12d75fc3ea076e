#!/bin/bash
# round 2, GPU call J: device-resident StochasticPart (drop-in test, K5 chain
# with device-resident fields against the host StochasticPart), the
# allocation-free scalar product alone in a fresh process and after the suite
mkdir -p gpurun_out
timeout 600 oracle/_ref/dropin_test > gpurun_out/r2j_dropin.log 2>&1
echo "dropin rc=$?" >> gpurun_out/r2j_dropin.log
for c in 128 256; do
  timeout 300 oracle/_ref/dropin_test chain $c 5 >> gpurun_out/r2j_chain.jsonl 2>> gpurun_out/r2j_chain.err
  timeout 300 oracle/_ref/dropin_test_host chain $c 5 >> gpurun_out/r2j_chain.jsonl 2>> gpurun_out/r2j_chain.err
done
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/r2j_noalloc_alone.log 2>&1
echo "alone rc=$?" >> gpurun_out/r2j_noalloc_alone.log
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r2j_gpu_tests.log 2>&1
echo "pytest rc=$?" >> gpurun_out/r2j_gpu_tests.log
tail -n 4 gpurun_out/r2j_dropin.log gpurun_out/r2j_noalloc_alone.log gpurun_out/r2j_gpu_tests.log
cat gpurun_out/r2j_chain.jsonl
