#!/bin/bash
# usage: tools/gpurun_retry.sh <timeout_s> <gpus> <command...>   -- retries while the pod answers "busy" (nothing charged)
to=$1; shift; gpus=$1; shift
for attempt in $(seq 1 40); do
  if [ "$gpus" -gt 1 ]; then /usr/local/graft/bin/gpurun --gpus $gpus --timeout $to -- "$@" > /tmp/gpurun_last.log 2>&1; rc=$?
  else /usr/local/graft/bin/gpurun --timeout $to -- "$@" > /tmp/gpurun_last.log 2>&1; rc=$?; fi
  if grep -q "status=transient" /tmp/gpurun_last.log || [ $rc -eq 3 ]; then echo "attempt $attempt: busy, retrying in 120 s"; sleep 120; continue; fi
  break
done
tail -200 /tmp/gpurun_last.log
exit $rc
