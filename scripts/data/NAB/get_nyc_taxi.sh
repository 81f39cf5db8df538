#!/bin/bash
# NYC taxi passengers series of the Numenta Anomaly Benchmark into <dir>/NAB/nyc_taxi (reference: scripts/data/NAB/nyc_taxi/
# get_nyc_taxi.sh).  Offline: a synthetic half-hourly series with the same columns (timestamp, value).
DIR="${1:-.}/NAB/nyc_taxi"
mkdir -p "$DIR"
if curl -fsS --max-time 20 -o "$DIR/nyc_taxi.csv" https://raw.githubusercontent.com/numenta/NAB/master/data/realKnownCause/nyc_taxi.csv 2>/dev/null; then
    echo "downloaded nyc_taxi.csv into $DIR"
else
    python "$(dirname "${BASH_SOURCE[0]}")/../gen_synthetic.py" nab "$DIR"
fi
