#!/bin/bash
# retry a gpurun command until it is accepted (status transient -> sleep and retry)
for i in $(seq 1 12); do
  out=$(/usr/local/graft/bin/gpurun --timeout ${GT:-900} -- "$1" 2>&1)
  echo "$out" | tail -${TAILN:-30}
  if ! echo "$out" | grep -q "status=transient"; then exit 0; fi
  echo "--- transient, retry $i"; sleep 90
done
