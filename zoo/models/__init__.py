"""Built-in models (reference: pyzoo/zoo/models, zoo/src/main/scala/.../models)."""
