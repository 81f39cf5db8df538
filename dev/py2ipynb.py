#!/usr/bin/env python
"""Generate the Jupyter notebooks under apps/ from the application scripts (the reference ships notebooks and converts them to
scripts for its CI, apps/ipynb2py.sh; here the scripts are the source and the notebooks are generated, so the two cannot drift).

    python dev/py2ipynb.py            # (re)write every notebook listed in NOTEBOOKS
    python dev/py2ipynb.py --check    # exit 1 if a committed notebook is stale

Cell rules: the module docstring becomes the title cell; a top-level comment block that starts with `# ##` starts a markdown cell;
blank-line separated top-level statement groups become code cells; the `sys.path` bootstrap of the scripts is replaced by one
cell that finds the repository root from the notebook's directory."""
import argparse
import ast
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
# notebook (relative to apps/) -> source script (relative to the repository root)
NOTEBOOKS = {
    "anomaly-detection/anomaly-detection-nyc-taxi.ipynb": "examples/anomalydetection/anomaly_detection.py",
    "anomaly-detection-hd/autoencoder-zoo.ipynb": "apps/anomaly-detection-hd/autoencoder_anomaly.py",
    "automl/nyc_taxi_dataset.ipynb": "examples/autots/autots_forecast.py",
    "dogs-vs-cats/transfer-learning.ipynb": "examples/pytorch/resnet_finetune_nnframes.py",
    "fraud-detection/fraud-detection.ipynb": "apps/fraud-detection/fraud_detection.py",
    "image-augmentation/image-augmentation.ipynb": "apps/image-augmentation/image_augmentation.py",
    "image-augmentation-3d/image-augmentation-3d.ipynb": "apps/image-augmentation-3d/image_augmentation_3d.py",
    "image-similarity/image-similarity.ipynb": "apps/image-similarity/image_similarity.py",
    "object-detection/object-detection.ipynb": "examples/objectdetection/ssd_finetune.py",
    "pytorch/face_generation.ipynb": "apps/pytorch/face_generation.py",
    "ray/parameter_server/sharded_parameter_server.ipynb": "examples/ray/parameter_server.py",
    "recommendation-ncf/ncf-explicit-feedback.ipynb": "examples/recommendation/ncf_explicit_feedback.py",
    "recommendation-wide-n-deep/wide_n_deep.ipynb": "examples/recommendation/wide_n_deep.py",
    "sentiment-analysis/sentiment.ipynb": "apps/sentiment-analysis/sentiment.py",
    "tfnet/image_classification_inference.ipynb": "apps/tfnet/image_classification_inference.py",
    "variational-autoencoder/using_variational_autoencoder_to_generate_digital_numbers.ipynb": "apps/variational-autoencoder/vae.py",
    "model-inference-examples/inference_model.ipynb": "examples/inference/inference_model.py",
    "model-inference-examples/cluster_serving_quick_start.ipynb": "examples/serving/quick_start.py",
}
BOOTSTRAP = ("import os\nimport sys\n\n_d = os.path.abspath(os.getcwd())\nwhile not os.path.exists(os.path.join(_d, 'zoo', '__init__.py')) and _d != '/':\n"
             "    _d = os.path.dirname(_d)\nsys.path.insert(0, _d)          # the repository root, wherever the notebook was started from\n"
             "sys.argv = sys.argv[:1]         # the scripts read options with argparse: use their defaults")


def _md(text):
    return {"cell_type": "markdown", "metadata": {}, "source": text.strip("\n").splitlines(keepends=True)}


def _code(text):
    return {"cell_type": "code", "execution_count": None, "metadata": {}, "outputs": [], "source": text.strip("\n").splitlines(keepends=True)}


def convert(src_path):
    with open(src_path) as fh:
        src = fh.read()
    tree = ast.parse(src)
    lines = src.splitlines()
    cells = []
    body = list(tree.body)
    doc = ast.get_docstring(tree)
    if doc:
        title = os.path.splitext(os.path.basename(src_path))[0].replace("_", " ")
        cells.append(_md(f"# {title}\n\n{doc}\n\n*Generated from `{os.path.relpath(src_path, ROOT)}` by `dev/py2ipynb.py`.*"))
        body = body[1:]
    cells.append(_code(BOOTSTRAP))
    # statement groups: consecutive top-level statements with no blank line between them
    groups, cur, prev_end = [], [], None
    for node in body:
        start = min([node.lineno] + [d.lineno for d in getattr(node, "decorator_list", [])])
        k = start - 2                                  # pull the comment lines directly above into the group
        while k >= 0 and lines[k].lstrip().startswith("#"):
            start = k + 1
            k -= 1
        if prev_end is not None and start - prev_end > 1 and cur:
            groups.append(cur)
            cur = []
        cur.append((start, node.end_lineno))
        prev_end = node.end_lineno
    if cur:
        groups.append(cur)
    for g in groups:
        text = "\n".join(lines[g[0][0] - 1: g[-1][1]])
        kept = []
        for ln in text.splitlines():
            s = ln.strip()
            if s.startswith("sys.path.insert(") or s in ("import _common  # noqa: F401", "import _common"):
                continue
            kept.append(ln)
        text = "\n".join(kept).strip("\n")
        if not text.strip() or all(l.strip() in ("import os", "import sys", "") for l in text.splitlines()):
            continue
        first = text.splitlines()[0]
        if first.startswith("# ##"):
            head = [first[2:]]
            rest = text.splitlines()[1:]
            while rest and rest[0].startswith("#"):
                head.append(rest.pop(0)[1:].strip())
            cells.append(_md("\n".join(head)))
            text = "\n".join(rest)
            if not text.strip():
                continue
        text = text.replace('if __name__ == "__main__":\n    main()', "main()")
        cells.append(_code(text))
    return {"cells": cells, "metadata": {"kernelspec": {"display_name": "Python 3", "language": "python", "name": "python3"},
                                         "language_info": {"name": "python"}}, "nbformat": 4, "nbformat_minor": 5}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--check", action="store_true")
    a = ap.parse_args()
    stale = []
    for nb, src in NOTEBOOKS.items():
        out = os.path.join(ROOT, "apps", nb)
        text = json.dumps(convert(os.path.join(ROOT, src)), indent=1) + "\n"
        if a.check:
            if not os.path.exists(out) or open(out).read() != text:
                stale.append(nb)
            continue
        os.makedirs(os.path.dirname(out), exist_ok=True)
        with open(out, "w") as fh:
            fh.write(text)
    if a.check and stale:
        print("stale notebooks:", stale)
        sys.exit(1)
    print(f"{len(NOTEBOOKS)} notebooks {'checked' if a.check else 'written'}")


if __name__ == "__main__":
    main()
