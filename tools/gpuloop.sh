#!/bin/bash
# gpuloop.sh <log> <gpurun args...>: retry gpurun while the pod answers busy (exit 3)
log=$1; shift
for i in $(seq 1 60); do
  /usr/local/graft/bin/gpurun "$@" > "$log" 2>&1
  rc=$?
  if [ $rc -ne 3 ]; then echo "gpurun exit $rc" >> "$log"; exit $rc; fi
  sleep 60
done
exit 3
