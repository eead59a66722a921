#!/bin/bash
# usage: tools/gpurun_retry.sh [gpurun options] -- 'command'   (retries while the pod answers busy)
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun "$@" > /tmp/gpurun_last.txt 2>&1
  rc=$?
  if grep -q "status=transient" /tmp/gpurun_last.txt || [ $rc -eq 3 ]; then sleep 90; continue; fi
  cat /tmp/gpurun_last.txt | tail -${TAIL:-40}
  exit $rc
done
echo "gave up after 40 tries"; exit 3
