"""TensorFlow graph export helpers (reference: pyzoo/zoo/util/tf.py:50-300: `export_tf` freezes a session graph into a folder
TFNet can load, `strip_unused`, `process_grad`).  No TensorFlow runtime exists here; the torch route is
`zoo.pipeline.inference.InferenceModel.load_torch` / `torch.jit.save`."""


def _needs_tf(name):
    raise NotImplementedError(f"zoo.util.tf.{name} freezes or edits a TensorFlow graph and needs the TensorFlow runtime, which is not part of "
                              "zoo-b200; export the model with torch.jit.save / torch.save and load it with InferenceModel.load_torch")


def process_grad(grad):
    _needs_tf("process_grad")


def export_tf(sess, folder, inputs, outputs, generate_backward=False, allow_non_differentiable_input=True):
    _needs_tf("export_tf")


def strip_unused(input_graph_def, input_tensor_names, output_tensor_names, placeholder_type_enum):
    _needs_tf("strip_unused")
