#!/bin/bash
# correctness of a library variant (op-level + generator parity) and same-box launch-list A/B against the baseline variant
mkdir -p gpurun_out
v=${1:-biasmma}
HFG_LIB=$PWD/build/variants/$v/libhifigan_b200.so timeout 900 python -m pytest tests/test_ops_gpu.py tests/test_generator_gpu.py -m gpu -x -q -k "resblock or golden or seeded or full_length or fused" > gpurun_out/pytest_$v.log 2>&1; echo "pytest($v) rc=$?"; tail -5 gpurun_out/pytest_$v.log
VARIANTS="rr $v" bash tools/r2_ab_variants.sh
