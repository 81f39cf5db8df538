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
    env = dict(os.environ, ZOO_FORCE_CPU="1", OMP_NUM_THREADS="2", EPOCHS="2", STEPS="40", ITERS="60", BAGS="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "examples", script), *args], capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    assert r.returncode == 0 and expect in r.stdout, r.stdout[-1500:] + r.stderr[-3000:]
