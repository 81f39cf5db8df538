#!/bin/bash
# Jupyter inside a hyperzoo pod (reference: docker/hyperzoo/start-notebook-k8s.sh / start-notebook-spark.sh).
exec bash "$(dirname "${BASH_SOURCE[0]}")/../zoo/start-notebook.sh" "$@"
