#!/bin/bash
# debug matrix for the TMA paths
cd "$(dirname "$0")/.."
for envs in "B200IMG_NO_TMA_LOAD=1" "B200IMG_NO_TMA_STORE=1" "B200IMG_DEBUG=1" "B200IMG_DEBUG=1 B200IMG_NO_TMA_STORE=1" "B200IMG_DEBUG=1 B200IMG_NO_TMA_LOAD=1"; do
  echo "== $envs"
  env $envs python scripts/dbg_cases.py 3 5 2>&1 | tail -2
done
