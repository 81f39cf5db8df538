#!/bin/bash
# Rebuild the native libraries, then run a command on a B200 through gpurun.
# usage: ./scripts_gpu.sh [--gpus N] [--timeout S] -- '<command>'
set -e
cd "$(dirname "$0")"
python -m zoo.native.build >/dev/null
exec /usr/local/graft/bin/gpurun "$@"
