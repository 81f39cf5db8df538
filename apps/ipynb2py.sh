#!/bin/bash
# Notebook -> script, the way the reference's apps/ipynb2py.sh does it (cell magics commented out, then nbconvert);
# without jupyter the code cells are extracted directly from the notebook's JSON.
#   usage: ./ipynb2py.sh <file-name without extension>
if [ $# -ne "1" ]; then
    echo "Usage: ./ipynb2py.sh <file-name without extension>"
    exit 1
fi
if command -v jupyter >/dev/null 2>&1; then
    cp "$1.ipynb" "$1.tmp.ipynb"
    sed -i 's/%%/#/; s/%pylab/#/' "$1.tmp.ipynb"
    jupyter nbconvert "$1.tmp.ipynb" --to python && mv "$1.tmp.py" "$1.py"
    rm -f "$1.tmp.ipynb"
else
    python - "$1" <<'PY'
import json, sys
nb = json.load(open(sys.argv[1] + ".ipynb"))
with open(sys.argv[1] + ".py", "w") as out:
    out.write("# -*- coding: utf-8 -*-\n")
    for cell in nb["cells"]:
        if cell["cell_type"] == "code":
            src = "".join(cell["source"])
            out.write("\n".join("#" + l if l.lstrip().startswith("%") else l for l in src.split("\n")) + "\n\n")
PY
fi
