#!/bin/bash
# e2e time of the one-call host path for a few pipeline schedules (tools/e2e_sched.py), after the parity tests of the extractor
python -m pytest tests -m gpu -x -q -k "extract or ref or golden or wire" 2>&1 | tail -2
bash tools/gpu_check.sh 2>&1 | tail -1
for s in "" "32,150,0,128,64" "32,150,0,160,64" "32,140,0,128,48" "24,150,0,128,64" "32,150,0,96,64" "48,150,0,128,64" "32,170,0,128,32" "32,150,0,128,96"; do
  if [ -z "$s" ]; then python tools/e2e_sched.py; else ORBFE_PIPE_SCHED=$s python tools/e2e_sched.py; fi
done
