#!/bin/bash
# round 2, GPU call Z (1 GPU): cache-hint experiment on the specialised passes (C3, 30 qubits)
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
for h in 0 1 2 3; do
  echo "== B200SV_JIT_MEMHINT=$h"
  B200SV_JIT_MEMHINT=$h timeout 300 python scripts/jit30.py 30 4:11:1:3 2>&1 | grep "rep [23]\|expval" | cut -c1-200
done
