# Source me: environment of a zoo-b200 checkout (reference: scripts/analytics-zoo-env.sh).
export ZOO_HOME="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"
export ANALYTICS_ZOO_HOME="${ANALYTICS_ZOO_HOME:-$ZOO_HOME}"      # the reference's name for the same directory
export PYTHONPATH="$ZOO_HOME:$PYTHONPATH"
export PATH="$ZOO_HOME/scripts:$ZOO_HOME/scripts/cluster-serving:$PATH"
# knobs the framework reads (zoo/common/nncontext.py): ZOO_NUM_MKLTHREADS / OMP_NUM_THREADS, ZOO_FORCE_CPU=1,
# ZOO_SYMM=vmm|ipc, ZOO_COMM_BACKEND=auto|p2p|nccl|gloo, ZOO_FAULT_INJECT=rank:iteration
