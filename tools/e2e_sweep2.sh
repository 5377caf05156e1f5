#!/bin/bash
# one-call e2e for a few pipeline schedules, then the stage timeline of the default
for s in "" "32,150,0,128,64" "32,150,0,128,32" "32,150,0,96,32" "32,140,0,96,48" "48,150,0,160,64" "32,130,0,64,32" "16,150,0,128,64" "32,150,0,128,128" "64,150,0,128,64"; do
  if [ -z "$s" ]; then python tools/e2e_onecall.py; else ORBFE_PIPE_SCHED=$s python tools/e2e_onecall.py; fi
done 2>&1 | grep one-call
ORBFE_PIPE_DUAL=0 python tools/e2e_onecall.py 2>&1 | grep one-call
REPS=1 ORBFE_PIPE_TRACE=1 python tools/e2e_onecall.py 2>&1 | tail -14
