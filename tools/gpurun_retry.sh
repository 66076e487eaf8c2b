#!/bin/bash
# usage: tools/gpurun_retry.sh <logfile> <gpurun args...>   -- retries while gpurun answers "no box / slot free" (exit 3)
log=$1; shift
for try in $(seq 1 40); do
  /usr/local/graft/bin/gpurun "$@" > "$log" 2>&1
  rc=$?
  if [ $rc -ne 3 ]; then echo "gpurun exit $rc after $try tries"; tail -25 "$log"; exit $rc; fi
  sleep 120
done
echo "gave up"; exit 3
