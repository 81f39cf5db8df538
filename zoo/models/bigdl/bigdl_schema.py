"""Field numbers of BigDL's model serialization (`bigdl.proto`, BigDL 0.5+: `Module.saveModule` / `ZooModel.saveModel`), written
down from the wire format of saved models (reference artefacts: zoo/src/test/resources/models/bigdl/bigdl_lenet.model,
models/zoo_keras/small_seq.model) for :mod:`zoo.util.protowire`.  Only the fields the loader and the writer use are named.

A saved model is one `BigDLModule`.  Parameters are `BigDLTensor`s whose `storage` carries only an id; the data of every storage
sits once in the ROOT module's attribute `global_storage` (a `NameAttrList`: tensor id -> AttrValue{tensorValue with data}).
"""

STORAGE = {1: ("datatype", "enum"), 2: ("float_data", "packed_float"), 3: ("double_data", "packed_double"), 6: ("int_data", "packed_varint"),
           7: ("long_data", "packed_varint"), 9: ("id", "varint")}
TENSOR = {1: ("datatype", "enum"), 2: ("size", "packed_varint"), 3: ("stride", "packed_varint"), 4: ("offset", "varint"),
          5: ("dimension", "varint"), 6: ("nElements", "varint"), 7: ("isScalar", "bool"), 8: ("storage", "msg", STORAGE), 9: ("id", "varint"),
          10: ("tensorType", "enum")}
SHAPE = {}
SHAPE.update({1: ("shapeType", "enum"), 2: ("ssize", "varint"), 3: ("shapeValue", "packed_varint"), 4: ("shape", "msg", SHAPE)})
INIT_METHOD = {1: ("methodType", "enum"), 2: ("data", "packed_double")}
MODULE, ATTR, ARRAY, NAME_ATTR_LIST, ATTR_ENTRY = {}, {}, {}, {}, {}
ARRAY.update({1: ("size", "varint"), 2: ("datatype", "enum"), 3: ("i32", "packed_varint"), 4: ("i64", "packed_varint"), 5: ("flt", "packed_float"),
              6: ("dbl", "packed_double"), 7: ("str", "string"), 8: ("boolean", "packed_varint"), 10: ("tensor", "msg", TENSOR),
              13: ("bigDLModule", "msg", MODULE), 14: ("nameAttrList", "msg", NAME_ATTR_LIST), 17: ("shape", "msg", SHAPE)})
ATTR.update({1: ("dataType", "enum"), 2: ("subType", "string"), 3: ("int32Value", "varint"), 4: ("int64Value", "varint"), 5: ("floatValue", "float"),
             6: ("doubleValue", "double"), 7: ("stringValue", "string"), 8: ("boolValue", "bool"), 10: ("tensorValue", "msg", TENSOR),
             12: ("initMethodValue", "msg", INIT_METHOD), 13: ("bigDLModuleValue", "msg", MODULE), 14: ("nameAttrListValue", "msg", NAME_ATTR_LIST),
             15: ("arrayValue", "msg", ARRAY), 16: ("dataFormatValue", "enum"), 18: ("shape", "msg", SHAPE)})
ATTR_ENTRY.update({1: ("key", "string"), 2: ("value", "msg", ATTR)})
NAME_ATTR_LIST.update({1: ("name", "string"), 2: ("attr", "msg", ATTR_ENTRY)})
MODULE.update({1: ("name", "string"), 2: ("subModules", "msg", MODULE), 3: ("weight", "msg", TENSOR), 4: ("bias", "msg", TENSOR),
               5: ("preModules", "string"), 6: ("nextModules", "string"), 7: ("moduleType", "string"), 8: ("attr", "msg", ATTR_ENTRY),
               9: ("version", "string"), 10: ("train", "bool"), 11: ("namePostfix", "string"), 12: ("id", "varint"),
               13: ("inputShape", "msg", SHAPE), 14: ("outputShape", "msg", SHAPE), 15: ("hasParameters", "bool"),
               16: ("parameters", "msg", TENSOR)})

# DataType enum values used by the writer
DT_INT32, DT_INT64, DT_FLOAT, DT_DOUBLE, DT_STRING, DT_BOOL = 0, 1, 2, 3, 4, 5
DT_TENSOR, DT_MODULE, DT_NAME_ATTR_LIST, DT_ARRAY, DT_DATA_FORMAT = 10, 13, 14, 15, 16
