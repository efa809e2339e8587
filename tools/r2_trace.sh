#!/bin/bash
# role timelines of the fused kernel from a trace build (HFG_BUILD_TRACE=1 rebuilds the library on the box, in place)
mkdir -p gpurun_out
HFG_BUILD_TRACE=1 python -c "from neuraltexttospeech_b200.build import build; build(force=True)" > gpurun_out/trace_build.log 2>&1
for spec in "3 1 1" "7 3 1" "11 5 1" "3 1 2"; do
  HFG_BUILD_TRACE=1 python tools/trace_resblock.py $spec > "gpurun_out/trace_rb_$(echo $spec | tr ' ' '_').txt" 2>&1
  tail -8 "gpurun_out/trace_rb_$(echo $spec | tr ' ' '_').txt"
done
