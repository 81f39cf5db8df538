#!/bin/bash
# GloVe 6B vectors into <dir>/glove.6B (reference: scripts/data/glove/get_glove.sh downloads glove.6B.zip).  Offline: a synthetic
# glove.6B.50d.txt in the same text format.
DIR="${1:-.}/glove.6B"
mkdir -p "$DIR"
if curl -fsS --max-time 20 -o "$DIR/glove.6B.zip" http://nlp.stanford.edu/data/glove.6B.zip 2>/dev/null && unzip -q -o "$DIR/glove.6B.zip" -d "$DIR"; then
    echo "downloaded GloVe into $DIR"
else
    python "$(dirname "${BASH_SOURCE[0]}")/../gen_synthetic.py" glove "$DIR"
fi
