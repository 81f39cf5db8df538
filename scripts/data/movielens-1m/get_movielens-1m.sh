#!/bin/bash
# MovieLens-1M into <dir>/movielens-1m (reference: scripts/data/movielens-1m/get_movielens-1m.sh downloads ml-1m.zip from
# files.grouplens.org).  With network access the same download is tried; otherwise a synthetic ratings.dat of the same shape.
DIR="${1:-.}/movielens-1m"
mkdir -p "$DIR"
if curl -fsS --max-time 20 -o "$DIR/ml-1m.zip" http://files.grouplens.org/datasets/movielens/ml-1m.zip 2>/dev/null && unzip -q -o "$DIR/ml-1m.zip" -d "$DIR"; then
    echo "downloaded MovieLens-1M into $DIR"
else
    python "$(dirname "${BASH_SOURCE[0]}")/../gen_synthetic.py" movielens "$DIR"
fi
