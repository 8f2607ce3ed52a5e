#!/bin/bash
# usage: tools/gpu/retry.sh <timeout> <script> ; retries while the pod answers busy (exit 3)
for i in $(seq 1 20); do
  /usr/local/graft/bin/gpurun --timeout $1 -- "bash $2" ; rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 120
done
exit 3
