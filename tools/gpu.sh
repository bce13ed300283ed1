#!/bin/bash
# build in-tree (nvcc cross-compiles here), then run a command on a B200 box:  tools/gpu.sh [--timeout S] -- '<cmd>'
set -e
cd "$(dirname "$0")/.."
python sai2-primitives-perso_b200/build.py > /dev/null
make -s -C oracle
exec /usr/local/graft/bin/gpurun "$@"
