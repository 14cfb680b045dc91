#!/bin/bash
# usage: scripts/gpucall.sh <logfile> <timeout> -- '<cmd>'   retries while the pod answers "transient / nothing was charged"
log=$1; shift; to=$1; shift
for attempt in 1 2 3 4 5 6 7 8 9 10 11 12; do
  /usr/local/graft/bin/gpurun --timeout $to "$@" > $log 2>&1
  rc=$?
  if grep -q "status=transient" $log; then sleep 120; continue; fi
  break
done
echo "rc=$rc attempts=$attempt" >> $log
