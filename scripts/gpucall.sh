#!/bin/bash
# usage: scripts/gpucall.sh <timeout_s> <logfile> [--gpus N] -- <command...>   (retries while the pod answers "busy", exit code 3)
T=$1; LOG=$2; shift 2
EXTRA=()
while [ "$1" != "--" ]; do EXTRA+=("$1"); shift; done
shift
for try in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout "$T" "${EXTRA[@]}" -- "$@" > "$LOG" 2>&1
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 90
done
exit 3
