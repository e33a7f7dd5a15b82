#!/bin/bash
# gpurun_retry.sh <log> <gpurun args...>: re-submit while the pod answers "busy / transient" (nothing charged)
LOG=$1; shift
for i in 1 2 3 4 5 6 7 8 9 10; do
  /usr/local/graft/bin/gpurun "$@" > "$LOG" 2>&1
  if grep -q "status=transient\|status=busy\|retry in a few minutes" "$LOG"; then sleep 60; continue; fi
  break
done
