#!/bin/bash
# HiCS synthetic multi-dimensional outlier data sets into <dir>/HiCS (reference: scripts/data/HiCS/get_HiCS.sh downloads
# synth_multidim_100_000.csv for the anomaly-detection-hd app).  Offline: data of the same form is generated -- clusters in a few
# correlated subspaces, the last column marks the injected outliers.
DIR="${1:-.}/HiCS"
mkdir -p "$DIR"
python "$(dirname "${BASH_SOURCE[0]}")/../gen_synthetic.py" hics "$DIR"
