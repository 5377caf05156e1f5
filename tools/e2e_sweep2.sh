#!/bin/bash
# one-call e2e for a few pipeline schedules (first,grow%,add,max,tail), then the stage timeline of the default
for s in "" "32,150,0,128,48" "32,150,0,128,32" "32,150,0,128,16" "32,150,0,112,32" "48,150,0,128,32" "32,200,0,128,32" "32,150,0,160,32"; do
  if [ -z "$s" ]; then python tools/e2e_onecall.py; else ORBFE_PIPE_SCHED=$s python tools/e2e_onecall.py; fi
done 2>&1 | grep one-call
REPS=1 ORBFE_PIPE_TRACE=1 python tools/e2e_onecall.py 2>&1 | tail -12
