#!/bin/bash
# Start Jupyter with zoo on the path (reference: scripts/jupyter-with-zoo.sh).
here="$(cd "$(dirname "$0")" && pwd)"
export PYTHONPATH="$(cd "$here/.." && pwd):$PYTHONPATH"
exec jupyter notebook --notebook-dir="${1:-.}" --ip=0.0.0.0 --no-browser --allow-root
