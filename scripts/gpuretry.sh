#!/bin/bash
# gpuretry.sh <log> <timeout-seconds> [--gpus N] -- '<command>': gpurun with retries while the pod answers busy/transient
# (exit 3 or a "transient" verdict: nothing was charged).  Appends DONE rc=<rc> to the log when the call has ended.
log=$1; shift
tmo=$1; shift
extra=()
while [ "$1" != "--" ]; do extra+=("$1"); shift; done
shift
for try in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout "$tmo" "${extra[@]}" -- "$1" > "$log" 2>&1
  rc=$?
  if [ $rc -eq 3 ] || grep -q '"status": *"transient"\|status=transient' "$log"; then
    echo "try $try: rc=$rc busy/transient, retrying in 90 s" >> "$log.tries"
    sleep 90
    continue
  fi
  break
done
echo "DONE rc=$rc" >> "$log"
