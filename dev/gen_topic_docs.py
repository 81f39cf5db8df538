#!/usr/bin/env python
"""Generate the per-topic reference pages of the docs site from the code: one page per topic of the reference's docs tree
(docs/docs/APIGuide/**, KerasStyleAPIGuide/**, Keras2StyleAPIGuide/**, Zouwu/API/**, ClusterServingGuide/APIGuide.md), each with the
constructor signature, the full docstring (which cites the reference file and lines the class restates) and a method table for
every public class / function of the topic.  Hand-written guides live next to these pages; these cannot drift from the code.

    python dev/gen_topic_docs.py
"""
import importlib
import inspect
import os
import sys
import textwrap

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

K = "zoo.pipeline.api.keras"
# page path -> (title, intro, [(module, None | [names] | predicate-on-name)])
PAGES = {
    "APIGuide/AutoML/recipe.md": ("AutoML: recipes", "A recipe defines the search space and the stopping rule of a hyper-parameter search.",
                                  [("zoo.automl.config.recipe", None)]),
    "APIGuide/AutoML/time-sequence-pipeline.md": ("AutoML: TimeSequencePipeline", "The fitted feature transformer + model a search returns.",
                                                  [("zoo.automl.pipeline.time_sequence", None)]),
    "APIGuide/AutoML/time-sequence-predictor.md": ("AutoML: TimeSequencePredictor", "Searches a forecasting pipeline for a time series.",
                                                   [("zoo.automl.regression.time_sequence_predictor", None), ("zoo.automl.feature.time_sequence", None)]),
    "APIGuide/FeatureEngineering/featureset.md": ("FeatureSet", "The training data container: memory tiers, sharding, batches.",
                                                  [("zoo.feature.common", ["FeatureSet", "MemoryType", "MiniBatch", "Sample"]), ("zoo.feature.pmem", None)]),
    "APIGuide/FeatureEngineering/image.md": ("Image feature engineering", "ImageSet, image preprocessing ops (2-D and 3-D), ROI ops, SequenceFiles of images.",
                                             [("zoo.feature.image.imageset", None), ("zoo.feature.image.imagePreprocessing", None),
                                                     ("zoo.feature.image.roi", None),
                                              ("zoo.feature.image.seqfile", None), ("zoo.feature.image3d.transformation", None)]),
    "APIGuide/FeatureEngineering/relation.md": ("Relations", "Pairs of text ids with a label, the input of ranking models.",
                                                [("zoo.feature.common", ["Relation", "Relations"])]),
    "APIGuide/FeatureEngineering/text.md": ("Text feature engineering", "TextSet, TextFeature and the text transformers.",
                                            [("zoo.feature.text.text_set", None), ("zoo.feature.text.text_feature", None),
                                                    ("zoo.feature.text.transformer", None)]),
    "APIGuide/FeatureEngineering/preprocessing.md": ("Preprocessing chains", "Composable record transformers and the batchers.",
                                                     [("zoo.feature.common", lambda n: n not in ("FeatureSet", "MemoryType", "MiniBatch", "Sample",
                                                             "Relation", "Relations"))]),
    "APIGuide/Models/anomaly-detection.md": ("AnomalyDetector", "", [("zoo.models.anomalydetection", None)]),
    "APIGuide/Models/image-classification.md": ("ImageClassifier", "", [("zoo.models.image.imageclassification", None), ("zoo.models.image.common", None)]),
    "APIGuide/Models/object-detection.md": ("ObjectDetector and SSD", "", [("zoo.models.image.objectdetection.object_detector", None),
            ("zoo.models.image.objectdetection.ssd", None),
                                                                            ("zoo.models.image.objectdetection.dataset", None),
                                                                                    ("zoo.models.image.objectdetection.evaluation", None)]),
    "APIGuide/Models/recommendation.md": ("Recommendation models", "", [("zoo.models.recommendation", None), ("zoo.models.recommendation.utils", None)]),
    "APIGuide/Models/seq2seq.md": ("Seq2seq", "", [("zoo.models.seq2seq", None)]),
    "APIGuide/Models/text-classification.md": ("TextClassifier", "", [("zoo.models.textclassification", None)]),
    "APIGuide/Models/text-matching.md": ("KNRM / TextMatcher", "", [("zoo.models.textmatching", None), ("zoo.models.common.ranker", None)]),
    "APIGuide/Models/model-files.md": ("Foreign model files", "BigDL / zoo saved models and Caffe models as torch modules.",
                                       [("zoo.models.bigdl", None), ("zoo.models.caffe", None), ("zoo.models.common.zoo_model", None)]),
    "APIGuide/PipelineAPI/estimator.md": ("Estimator", "", [("zoo.pipeline.estimator.estimator", None), ("zoo.pipeline.estimator.local_estimator", None),
                                                             ("zoo.common.triggers", None), ("zoo.pipeline.straggler", None)]),
    "APIGuide/PipelineAPI/inference.md": ("InferenceModel", "", [("zoo.pipeline.inference.inference_model", None),
            ("zoo.pipeline.inference.abstract_inference_model", None)]),
    "APIGuide/PipelineAPI/math.md": ("autograd math", "Functions over `Variable`s.", [("zoo.pipeline.api.autograd", lambda n: n[0].islower())]),
    "APIGuide/PipelineAPI/variable.md": ("autograd Variable, Parameter, Lambda, CustomLoss", "", [("zoo.pipeline.api.autograd", lambda n: n[0].isupper())]),
    "APIGuide/PipelineAPI/net.md": ("Net, TorchNet, TorchCriterion, TFNet, GraphNet", "", [("zoo.pipeline.api.net.net_load", None),
            ("zoo.pipeline.api.net.torch_net", None),
                                                                                             ("zoo.pipeline.api.net.torch_criterion", None),
                                                                                                     ("zoo.pipeline.api.net.tfnet", None),
                                                                                             ("zoo.pipeline.api.net.graph_net", None),
                                                                                                     ("zoo.pipeline.api.onnx.onnx_loader", None)]),
    "APIGuide/PipelineAPI/nnframes.md": ("NNFrames", "", [("zoo.pipeline.nnframes.nn_classifier", None), ("zoo.pipeline.nnframes.nn_image_reader", None)]),
    "APIGuide/TFPark/bert-classifier.md": ("BERT estimators", "", [("zoo.tfpark.text.estimator", None)]),
    "APIGuide/TFPark/estimator.md": ("TFEstimator", "", [("zoo.tfpark.estimator", None)]),
    "APIGuide/TFPark/gan-estimator.md": ("GANEstimator", "", [("zoo.tfpark.gan.gan_estimator", None)]),
    "APIGuide/TFPark/model.md": ("KerasModel", "", [("zoo.tfpark.model", None)]),
    "APIGuide/TFPark/text-models.md": ("Text models", "", [("zoo.tfpark.text.keras", None)]),
    "APIGuide/TFPark/tf-dataset.md": ("TFDataset", "", [("zoo.tfpark.tf_dataset", None), ("zoo.tfpark.tfrecord", None)]),
    "APIGuide/TFPark/tf-optimizer.md": ("TFOptimizer", "", [("zoo.tfpark.tf_optimizer", None), ("zoo.tfpark.zoo_optimizer", None)]),
    "APIGuide/TFPark/tf-pred.md": ("TFPredictor and TFNet", "", [("zoo.tfpark.tf_predictor", None), ("zoo.tfpark.graphdef", ["GraphDefModule",
            "read_graphdef", "write_graphdef"]),
                                                                 ("zoo.tfpark.saved_model", None)]),
    "KerasStyleAPIGuide/Layers/core.md": ("Core layers", "", [(K + ".layers.core", None), (K + ".layers.torch", None)]),
    "KerasStyleAPIGuide/Layers/convolutional.md": ("Convolutional layers", "", [(K + ".layers.convolutional", None), (K + ".layers.local", None)]),
    "KerasStyleAPIGuide/Layers/pooling.md": ("Pooling layers", "", [(K + ".layers.pooling", None)]),
    "KerasStyleAPIGuide/Layers/recurrent.md": ("Recurrent layers", "", [(K + ".layers.recurrent", None), (K + ".layers.convolutional_recurrent", None)]),
    "KerasStyleAPIGuide/Layers/embedding.md": ("Embedding layers", "", [(K + ".layers.embeddings", None)]),
    "KerasStyleAPIGuide/Layers/normalization.md": ("Normalization layers", "", [(K + ".layers.normalization", None)]),
    "KerasStyleAPIGuide/Layers/dropout.md": ("Dropout and noise layers", "", [(K + ".layers.noise", None), (K + ".layers.core", lambda n: "Dropout" in n)]),
    "KerasStyleAPIGuide/Layers/activation.md": ("Activations", "", [(K + ".layers.core", ["Activation"]), (K + ".layers.base",
            lambda n: "activation" in n.lower())]),
    "KerasStyleAPIGuide/Layers/advanced-activation.md": ("Advanced activations", "", [(K + ".layers.advanced_activations", None)]),
    "KerasStyleAPIGuide/Layers/self-attention.md": ("TransformerLayer and BERT", "", [(K + ".layers.self_attention", None)]),
    "KerasStyleAPIGuide/Layers/wrappers.md": ("Wrappers", "", [(K + ".layers.wrappers", None)]),
    "KerasStyleAPIGuide/Layers/initialization.md": ("Initializers and regularizers", "", [(K + ".regularizers", None), (K + ".layers.base",
            lambda n: "init" in n.lower())]),
    "KerasStyleAPIGuide/Optimization/objectives.md": ("Objectives (losses)", "", [(K + ".objectives", None)]),
    "KerasStyleAPIGuide/Optimization/optimizers.md": ("Optimizers and learning-rate schedules", "", [(K + ".optimizers", None)]),
    "KerasStyleAPIGuide/Optimization/training.md": ("compile / fit / evaluate / predict", "", [(K + ".engine.topology", None), (K + ".metrics", None)]),
    "Keras2StyleAPIGuide/Layers/layers.md": ("Keras-2 style layers", "The Keras-2 argument names over the same layers.",
            [("zoo.pipeline.api.keras2.layers", None)]),
    "Zouwu/API/AutoTSTrainer.md": ("AutoTSTrainer", "", [("zoo.zouwu.autots.forecast", ["AutoTSTrainer"])]),
    "Zouwu/API/TSPipeline.md": ("TSPipeline", "", [("zoo.zouwu.autots.forecast", ["TSPipeline"])]),
    "Zouwu/API/LSTMForecaster.md": ("LSTMForecaster", "", [("zoo.zouwu.model.forecast", ["LSTMForecaster", "Forecaster"])]),
    "Zouwu/API/MTNetForecaster.md": ("MTNetForecaster", "", [("zoo.zouwu.model.forecast", ["MTNetForecaster"])]),
    "ClusterServingGuide/APIGuide.md": ("Cluster Serving API", "Client queues and the server object.", [("zoo.serving.client", None),
            ("zoo.serving.server", None)]),
    "APIGuide/Runtime/context.md": ("Contexts and launchers", "", [("zoo.common.nncontext", None), ("zoo.util.launcher", None), ("zoo.util.spark", None)]),
    "APIGuide/Runtime/ray.md": ("RayContext and trainers on it", "", [("zoo.ray.raycontext", None), ("zoo.ray.mxnet.mxnet_trainer", None),
            ("zoo.orca.learn.pytorch.estimator", None)]),
    "APIGuide/Runtime/xshards.md": ("XShards", "", [("zoo.xshard.shard", None), ("zoo.xshard.pandas.preprocessing", None)]),
    "APIGuide/Runtime/utilities.md": ("File systems, TensorBoard, dictionaries", "", [("zoo.util.filesystem", None), ("zoo.tensorboard.writer", None),
            ("zoo.common.dictionary", None),
                                                                                      ("zoo.common.tensor_operation", None), ("zoo.common.utils", None)]),
}


# what the recurring constructor arguments of the Keras-style layers mean (the reference documents them per layer)
ARG_DOC = {
    "input_shape": "shape of one input record without the batch dimension; needed on the first layer of a `Sequential` only",
    "name": "layer name (defaults to the lower-cased class name plus a counter)",
            "bigdl_type": "kept for signature compatibility; parameters are fp32 with bf16 shadows on the GPU",
    "dim_ordering": "`\"th\"` = channels first (the default, the layout the kernels use), `\"tf\"` = channels last",
            "data_format": "Keras-2 name of `dim_ordering`: `\"channels_first\"` / `\"channels_last\"`",
    "activation": "name of the activation applied to the output (`relu`, `tanh`, `sigmoid`, `softmax`, `linear`, ...) or None",
    "inner_activation": "activation of the recurrent gates (default `hard_sigmoid`, which the persistent RNN kernels implement natively)",
    "W_regularizer": "regularizer of the input weights, e.g. `l2(1e-4)`; its penalty is added to the training loss",
            "U_regularizer": "regularizer of the recurrent weights",
    "b_regularizer": "regularizer of the bias", "kernel_regularizer": "Keras-2 name of `W_regularizer`", "bias_regularizer": "Keras-2 name of `b_regularizer`",
    "border_mode": "`\"valid\"` (no padding) or `\"same\"` (output has the input's spatial size at stride 1)", "padding": "Keras-2 name of `border_mode`",
    "bias": "whether the layer has a bias term", "use_bias": "Keras-2 name of `bias`",
            "init": "weight initializer name (`glorot_uniform`, `he_normal`, `uniform`, `zero`, ...)",
    "kernel_initializer": "Keras-2 name of `init`", "bias_initializer": "initializer of the bias (Keras-2 style layers)",
    "nb_filter": "number of output channels", "filters": "Keras-2 name of `nb_filter`", "nb_row": "kernel height", "nb_col": "kernel width",
            "kernel_size": "kernel size (Keras-2 style)",
    "filter_length": "kernel length of a 1-D convolution", "kernel_dim1": "kernel depth of a 3-D convolution",
            "kernel_dim2": "kernel height of a 3-D convolution", "kernel_dim3": "kernel width of a 3-D convolution",
    "subsample": "strides of the convolution", "subsample_length": "stride of a 1-D convolution",
            "strides": "strides (pooling layers: defaults to `pool_size`)", "stride": "stride of a 1-D pooling layer",
    "atrous_rate": "dilation of the kernel", "pool_size": "pooling window", "pool_length": "pooling window of a 1-D pooling layer",
            "output_dim": "size of the output (units / hidden size / embedding width)",
    "units": "Keras-2 name of `output_dim`", "input_dim": "size of the input (vocabulary size for embeddings)", "input_length": "length of the input sequences",
    "return_sequences": "return the output of every time step instead of the last one", "go_backwards": "process the sequence in reverse",
    "p": "dropout probability", "rate": "Keras-2 name of the dropout probability", "alpha": "slope / scale parameter of the activation",
            "size": "up-sampling factor", "cropping": "how much to cut from each border",
    "mode": "merge mode (`sum`, `mul`, `concat`, `ave`, `cos`, `dot`, `max`, `sub`, `div`, `min`)",
            "concat_axis": "axis of the concatenation in `concat` mode", "layers": "the layers whose outputs are merged",
    "merge_mode": "how a `Bidirectional` wrapper combines the two directions", "layer": "the wrapped layer", "n": "repetition count",
            "dims": "permutation of the non-batch dimensions (1-based)",
    "target_shape": "output shape without the batch dimension (one `-1` is inferred)", "mask_value": "time steps equal to this value are skipped",
            "weights": "initial weights (list of arrays)",
    "trainable": "whether the weights receive gradients", "mask_zero": "treat id 0 as padding (its row stays zero)",
            "zero_based_id": "ids start at 0 (default) instead of 1 as in BigDL's LookupTable",
    "padding_value": "id whose embedding stays zero", "combiner": "`sum`, `mean` or `sqrtn` over the ids of a bag (SparseEmbedding)",
            "max_norm": "rows are renormalised to at most this L2 norm",
    "embedding_file": "GloVe-format text file (`word v1 v2 ...`)", "word_index": "word -> id map (ids start at 1; 0 is padding)",
            "epsilon": "small constant added to the variance", "momentum": "momentum of the moving statistics",
    "beta_init": "initializer of the shift", "gamma_init": "initializer of the scale", "hidden_size": "model width of the Transformer / BERT block",
            "n_block": "number of Transformer blocks",
    "n_head": "attention heads", "intermediate_size": "width of the position-wise feed-forward layer", "hidden_drop": "dropout on the block outputs",
            "attn_drop": "dropout on the attention probabilities",
    "initializer_range": "standard deviation of the truncated-normal initializer",
            "output_all_block": "return the output of every block instead of the last", "embedding_layer": "custom embedding layer for the inputs",
    "bidirectional": "attend in both directions (False = causal mask)", "vocab": "vocabulary size", "seq_len": "maximum sequence length",
            "nb_kernel": "kernel size (ConvLSTM; square kernels)",
    "depth_multiplier": "depth-wise output channels per input channel", "depthwise_regularizer": "regularizer of the depth-wise kernel",
            "pointwise_regularizer": "regularizer of the 1x1 kernel",
    "nb_feature": "number of linear pieces of a MaxoutDense", "sigma": "standard deviation of the noise", "theta": "threshold",
            "value": "the constant", "constant": "the constant", "axis": "axis the op acts on",
    "dim": "dimension the op acts on (1-based, batch excluded where the reference does so)", "dimension": "dimension the op acts on",
            "index": "index to select", "length": "length of the slice", "offset": "start of the slice",
    "power": "exponent", "scale": "multiplier", "shift": "additive term", "min_value": "lower clamp", "max_value": "upper clamp",
            "torch_layer": "a torch module to wrap as a zoo layer",
}


def clean_doc(doc):
    return textwrap.dedent(doc).strip() if doc else ""


def sig(obj):
    try:
        return str(inspect.signature(obj))
    except (TypeError, ValueError):
        return "(...)"


def first_sentence(doc):
    if not doc:
        return ""
    text = " ".join(doc.strip().split("\n\n")[0].split())
    cut = text.find(". ")
    return (text[:cut + 1] if 0 < cut < 240 else text[:240]).replace("|", "\\|")


def select(mod, rule):
    names = getattr(mod, "__all__", None) or [n for n in vars(mod) if not n.startswith("_")]
    out = []
    for n in names:
        o = getattr(mod, n, None)
        if o is None or inspect.ismodule(o) or not (inspect.isclass(o) or inspect.isfunction(o)):
            continue
        if not (getattr(o, "__module__", "") or "").startswith("zoo"):
            continue
        if rule is None:
            if getattr(o, "__module__", None) != mod.__name__ and not mod.__name__.endswith(("recommendation", "anomalydetection", "textclassification",
                                                                                             "textmatching", "seq2seq", "imageclassification",
                                                                                                     "common", "bigdl", "caffe", "layers")):
                continue                                        # re-exports are documented where they are defined
        elif callable(rule):
            if not rule(n):
                continue
        elif n not in rule:
            continue
        out.append((n, o))
    key = (lambda p: rule.index(p[0])) if isinstance(rule, list) else (lambda p: (not inspect.isclass(p[1]), p[0].lower()))
    return sorted(out, key=key)


def describe(name, obj):
    lines = []
    if inspect.isclass(obj):
        lines += [f"## {name}", "", "```python", f"{name}{sig(obj)}", "```", ""]
        doc = clean_doc(obj.__doc__) or clean_doc(getattr(obj.__init__, "__doc__", None) if obj.__init__ is not object.__init__ else None)
        if doc:
            lines += [doc, ""]
        if ".keras" in (obj.__module__ or ""):                  # Keras-style layers: explain the constructor arguments from the glossary
            try:
                params = [p for p in inspect.signature(obj).parameters if p in ARG_DOC]
            except (TypeError, ValueError):
                params = []
            if params:
                lines += ["Arguments:", ""] + [f"* `{p}`: {ARG_DOC[p]}" for p in params] + [""]
        rows = []
        for m, v in vars(obj).items():
            if m.startswith("_"):
                continue
            f = v.__func__ if isinstance(v, (staticmethod, classmethod)) else v
            if not inspect.isfunction(f):
                continue
            kind = "static " if isinstance(v, staticmethod) else ("class " if isinstance(v, classmethod) else "")
            s = sig(f)
            rows.append(f"| `{kind}{m}{s if len(s) < 110 else s[:107] + '...'}` | {first_sentence(f.__doc__)} |".replace("\n", " "))
        if rows:
            lines += ["| method | what |", "|---|---|"] + rows + [""]
        bases = [b.__name__ for b in obj.__mro__[1:] if b.__module__.startswith("zoo")]
        if bases:
            lines += ["Inherits: " + ", ".join(f"`{b}`" for b in bases[:4]), ""]
    else:
        lines += [f"## {name}", "", "```python", f"{name}{sig(obj)}", "```", ""]
        doc = clean_doc(obj.__doc__)
        if doc:
            lines += [doc, ""]
    return lines


def main():
    pages = names = 0
    for path, (title, intro, sources) in PAGES.items():
        out = [f"# {title}", "",
                f"*Generated from the code by `python dev/gen_topic_docs.py`; the docstrings cite the reference file and lines each item restates.*", ""]
        if intro:
            out += [intro, ""]
        seen = set()
        for modname, rule in sources:
            try:
                mod = importlib.import_module(modname)
            except Exception as e:
                out += [f"*`{modname}` failed to import: {type(e).__name__}: {e}*", ""]
                continue
            items = [(n, o) for n, o in select(mod, rule) if id(o) not in seen]
            if not items:
                continue
            head = clean_doc(mod.__doc__)
            out += [f"---", f"Module `{modname}`", ""]
            if head and rule is None:
                out += [head, ""]
            for n, o in items:
                seen.add(id(o))
                out += describe(n, o)
                names += 1
        full = os.path.join(ROOT, "docs", path)
        os.makedirs(os.path.dirname(full), exist_ok=True)
        with open(full, "w") as fh:
            fh.write("\n".join(out).rstrip() + "\n")
        pages += 1
    print(f"{pages} pages, {names} documented classes / functions")


if __name__ == "__main__":
    main()
