#!/bin/bash
# Reference: scripts/cluster-serving/cluster-with-numactl.sh starts one Spark worker per NUMA node (and checks hyper-threading).
# Here: one serving process per GPU, each bound to the CPUs / memory of the NUMA node its GPU hangs off, each with its own RESP port
# (base port + GPU index).  usage: cluster-with-numactl.sh [config.yaml] [base_port]
conf="${1:-config.yaml}"; base="${2:-6380}"
here="$(cd "$(dirname "$0")" && pwd)"; root="$(cd "$here/../.." && pwd)"
[ -f "$conf" ] || { echo "$conf not found: run cluster-serving-init first"; exit 1; }
n=$(nvidia-smi -L 2>/dev/null | wc -l); [ "$n" -gt 0 ] || { echo "no GPU visible"; exit 1; }
for g in $(seq 0 $((n - 1))); do
  node=$(nvidia-smi topo -m 2>/dev/null | awk -v g="GPU$g" '$1 == g {print $(NF)}' | grep -E '^[0-9]+$' || true)
  bind=""; if command -v numactl >/dev/null 2>&1 && [ -n "$node" ]; then bind="numactl --cpunodebind=$node --membind=$node"; fi
  CUDA_VISIBLE_DEVICES=$g PYTHONPATH="$root:$PYTHONPATH" nohup $bind python -m zoo.serving.server -c "$conf" --port $((base + g)) \
      > "cluster-serving.gpu$g.out" 2>&1 &
  echo $! > "cluster-serving.gpu$g.pid"
  echo "GPU $g: serving process $! (NUMA node ${node:-n/a}, port $((base + g)))"
done
touch running
