#!/bin/bash
# usage (pod args): submit.sh script.py [args...]      env: ZOO_GPUS (default: all visible), MASTER_PORT, ZOO_WATCHDOG_TIMEOUT
set -e
n="${ZOO_GPUS:-$(nvidia-smi -L | wc -l)}"
exec python -m torch.distributed.run --nnodes=1 --nproc-per-node "$n" --master-addr 127.0.0.1 --master-port "${MASTER_PORT:-29500}" "$@"
