#!/bin/bash
# usage: tools/gpu.sh <timeout-seconds> '<command>'  — rebuild the library first so the box never runs a stale .so
set -e
cd "$(dirname "$0")/.."
python -m concrete_b200.build > /tmp/build.log 2>&1 || { tail -30 /tmp/build.log; exit 1; }
make -C oracle liboracle.so > /dev/null 2>&1 || true
exec /usr/local/graft/bin/gpurun --timeout "$1" -- "$2"
