#!/usr/bin/env bash
# Runs one gpurun call, retrying while the pod answers "busy" (exit 3: nothing charged).
# Usage: profiles/gpucall.sh LOGFILE TIMEOUT_S [--gpus N] -- 'command'
LOG="$1"; TMO="$2"; shift 2
for attempt in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout "$TMO" "$@" > "$LOG" 2>&1
  rc=$?
  if [ $rc -ne 3 ]; then echo "gpucall rc=$rc attempt=$attempt" >> "$LOG"; exit $rc; fi
  sleep 90
done
echo "gpucall gave up" >> "$LOG"; exit 3
