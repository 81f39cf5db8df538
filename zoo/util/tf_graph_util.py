"""GraphDef utilities copied by the reference from TensorFlow (pyzoo/zoo/util/tf_graph_util.py).  They operate on TF GraphDefs;
without a TensorFlow runtime they raise, naming the torch route."""
from .tf import _needs_tf


def must_run_on_cpu(node, pin_variables_on_cpu=False):
    _needs_tf("must_run_on_cpu")


def extract_sub_graph(graph_def, dest_nodes):
    _needs_tf("extract_sub_graph")


def tensor_shape_from_node_def_name(graph, input_name):
    _needs_tf("tensor_shape_from_node_def_name")


def convert_variables_to_constants(sess, input_graph_def, output_node_names, variable_names_whitelist=None, variable_names_blacklist=None):
    _needs_tf("convert_variables_to_constants")


def remove_training_nodes(input_graph, protected_nodes=None):
    _needs_tf("remove_training_nodes")
