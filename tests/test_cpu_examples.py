"""Run a few of the example scripts end to end with small settings (the reference converts notebooks to scripts and
runs them with fewer epochs: apps/run-app-tests.sh)."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CASES = [
    ("recommendation/ncf_explicit_feedback.py", ["--epochs", "1", "--ratings", "20000", "--batch-size", "2000"], "validation"),
    ("recommendation/wide_n_deep.py", ["--epochs", "1", "--rows", "4096", "--batch-size", "512"], "Top1Accuracy"),
    ("keras/functional_api.py", [], "mse after"),
    ("serving/quick_start.py", [], "results dequeued"),
    ("inference/inference_model.py", [], "all equal: True"),
    ("qaranker/qa_ranker.py", [], "NDCG@3"),
    ("attention/transformer_sentiment.py", [], "accuracy"),
    ("autograd/custom_loss.py", [], "mae before"),
    ("tfpark/gan_train.py", [], "generated radius"),
    ("tfpark/bert_classifier.py", [], "probabilities of 4 records"),
    ("ray/parameter_server.py", [], "weights after training: [1.0, 2.0, 3.0, 4.0]"),
    ("ray/async_parameter_server.py", [], "updates applied: 180"),
    ("streaming/text_classification_stream.py", [], "200 lines classified"),
    ("imageclassification/predict.py", [], "predicted 6 images"),
    ("localestimator/lenet_local.py", [], "Top1Accuracy"),
    ("inception/train_imagenet_style.py", ["--epochs", "1"], "validation:"),
    ("tensorflow/tfpark/tf_optimizer_train.py", [], "held-out accuracy"),
    ("tensorflow/tfpark/estimator_dataset.py", [], "predicted 600 records"),
    ("tensorflow/tfpark/keras_dataset.py", [], "held-out loss"),
    ("tensorflow/tfnet/predict.py", [], "predicted 6 images"),
    ("tensorflow/freeze/freeze_graph.py", [], "matches the reference computation: True"),
    ("vnni/fp8_perf.py", [], "top-1 agreement with fp32"),
    ("openvino/predict.py", [], "predicted 8 images"),
    ("nnframes/image_inference.py", [], "images classified: 8 distinct classes: [1, 2]"),
    ("nnframes/image_transfer_learning.py", [], "validation accuracy"),
    ("chatbot/train_chatbot.py", [], "exact answers on 50 training queries"),
    ("resnet/train_imagenet.py", ["--images", "128"], "validation:"),
    ("ray/mxnet_lenet.py", ["-e", "1"], "held-out accuracy after 1 epochs on 2 workers"),
    ("ray/rl_pong.py", ["--iterations", "30"], "running mean reward"),
    ("ray/multiagent_two_trainers.py", ["--iterations", "2"], "episode reward (max 200)"),
    ("objectdetection/predict.py", ["--images", "2"], "in 2 images; visualizations written"),
    ("streaming/object_detection_stream.py", [], "stream ended: 4 images in 2 micro-batches, 4 result files"),
    ("pytorch/simple_training.py", [], "predictions match labels: True"),
    ("textclassification/news20.py", [], "classes: 5, vocabulary of the embedding file: 20000"),
    ("autograd/custom.py", [], "loss points written to TensorBoard"),
    ("tensorflow/tfpark/tf_optimizer_evaluate.py", [], "records in batches of 20: accuracy"),
    ("tensorflow/tfpark/keras_ndarray.py", [], "weights saved: True"),
    ("pytorch/inference_predict.py", [], "predicted classes of 10 images"),
    ("recommendation/ml1m_wide_n_deep.py", ["--epochs", "2"], "finished..."),
    ("../apps/tfnet/image_classification_inference.py", [], "classified 4 images"),
    ("../apps/model-inference-examples/text_classification_inference/simple_driver.py", [], "The predict class is:0"),
    ("../apps/model-inference-examples/recommendation_inference/simple_driver.py", [], "pairs whose predicted class is the generating one"),
    ("../apps/model-inference-examples/model_inference_streaming/text_classification_streaming.py", [], "40 lines classified"),
    ("../apps/fraud-detection/fraud_detection.py", [], "AUPRC"),
    ("../apps/sentiment-analysis/sentiment.py", [], "held-out loss"),
    ("../apps/variational-autoencoder/vae.py", [], "negative ELBO"),
    ("../apps/image-similarity/image_similarity.py", [], "precision@5"),
    ("../apps/image-augmentation/image_augmentation.py", [], "wrote 15 images"),
    ("../apps/image-augmentation-3d/image_augmentation_3d.py", [], "affine (scale)"),
    ("../apps/anomaly-detection-hd/autoencoder_anomaly.py", [], "injected outliers found"),
    ("../apps/pytorch/face_generation.py", [], "generated (16, 3, 32, 32)"),
]


@pytest.mark.parametrize("script,args,expect", CASES)
def test_example_runs(script, args, expect):
    env = dict(os.environ, OMP_NUM_THREADS="2", EPOCHS="2", STEPS="40", ITERS="60", BAGS="1")
    if os.environ.get("ZOO_EXAMPLES_DEVICE", "cpu") != "gpu":      # ZOO_EXAMPLES_DEVICE=gpu runs the same list on the visible GPU (dev check)
        env["ZOO_FORCE_CPU"] = "1"
    r = subprocess.run([sys.executable, os.path.join(ROOT, "examples", script), *args], capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    assert r.returncode == 0 and expect in r.stdout, r.stdout[-1500:] + r.stderr[-3000:]


def test_notebooks_are_generated_from_the_scripts_and_run():
    """apps/*.ipynb are generated by dev/py2ipynb.py; they must be current, valid nbformat-4 JSON, and their code must run."""
    import json
    import tempfile

    r = subprocess.run([sys.executable, os.path.join(ROOT, "dev", "py2ipynb.py"), "--check"], capture_output=True, text=True, cwd=ROOT)
    assert r.returncode == 0, r.stdout + r.stderr
    nb = os.path.join(ROOT, "apps", "tfnet", "image_classification_inference.ipynb")
    doc = json.load(open(nb))
    assert doc["nbformat"] == 4 and doc["cells"][0]["cell_type"] == "markdown"
    code = "\n\n".join("".join(c["source"]) for c in doc["cells"] if c["cell_type"] == "code")
    with tempfile.NamedTemporaryFile("w", suffix=".py", delete=False, dir=os.path.dirname(nb)) as fh:
        fh.write(code)
    try:
        env = dict(os.environ, ZOO_FORCE_CPU="1", OMP_NUM_THREADS="2")
        r = subprocess.run([sys.executable, fh.name], capture_output=True, text=True, timeout=600, env=env, cwd=os.path.dirname(nb))
    finally:
        os.unlink(fh.name)
    assert r.returncode == 0 and "classified 4 images" in r.stdout, r.stdout[-1000:] + r.stderr[-2000:]
