"""A custom layer and a custom loss in the functional API (reference example: pyzoo/zoo/examples/autograd/custom.py: `Input` ->
`Dense(1)` -> `Lambda(add_one)` wrapped in `Model(input, output)`, compiled with a Python function as the loss, TensorBoard
summaries, `get_weights`, `predict_local`)."""
import os
import sys
import tempfile

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _common  # noqa: F401
import numpy as np

from zoo.common.nncontext import init_nncontext
from zoo.pipeline.api.autograd import Lambda, abs, mean
from zoo.pipeline.api.keras.layers import Dense, Input
from zoo.pipeline.api.keras.models import Model
from zoo.pipeline.api.keras.optimizers import SGD


def mean_absolute_error(y_true, y_pred):
    return mean(abs(y_true - y_pred), axis=1)


def add_one_func(x):
    return x + 1.0


if __name__ == "__main__":
    init_nncontext("custom example")
    rng = np.random.RandomState(0)
    data_len = 1000
    X_ = rng.uniform(0, 1, (data_len, 2)).astype(np.float32)
    Y_ = ((2 * X_).sum(1) + 0.4).reshape([data_len, 1]).astype(np.float32)
    a = Input(shape=(2,))
    b = Dense(1)(a)
    c = Lambda(function=add_one_func)(b)
    model = Model(input=a, output=c)
    model.compile(optimizer=SGD(learningrate=1e-2), loss=mean_absolute_error)
    log_dir = tempfile.mkdtemp(prefix="zoo_custom_")
    model.set_tensorboard(log_dir, "customized layer and loss")
    model.fit(x=X_, y=Y_, batch_size=32, nb_epoch=int(os.environ.get("CUSTOM_EPOCHS", 60)), distributed=False)
    w = model.get_weights()
    pred = model.predict_local(X_) if hasattr(model, "predict_local") else model.predict(X_)
    print("weights", np.round(np.asarray(w[0]).reshape(-1), 2).tolist(), "bias", np.round(np.asarray(w[1]).reshape(-1), 2).tolist(),
            "(target 2, 2 and 0.4 - 1)")
    print("mean absolute error on the training data: %.4f; loss points written to TensorBoard: %d" %
          (float(np.abs(np.asarray(pred) - Y_).mean()), len(model.get_train_summary("Loss"))))
