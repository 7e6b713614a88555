#!/bin/bash
# usage: tools/gpurun_retry.sh <gpus> <timeout> <command...>   — retries while the pod answers "transient" (no slot, nothing charged)
gpus=$1; shift; tmo=$1; shift
for i in $(seq 1 30); do
  if [ "$gpus" = "1" ]; then out=$(/usr/local/graft/bin/gpurun --timeout $tmo -- "$@" 2>&1); else out=$(/usr/local/graft/bin/gpurun --gpus $gpus --timeout $tmo -- "$@" 2>&1); fi
  if echo "$out" | grep -q "status=transient\|backing off\|another call"; then sleep 90; continue; fi
  echo "$out" | tail -40; exit 0
done
echo "gave up: $out" | tail -5
