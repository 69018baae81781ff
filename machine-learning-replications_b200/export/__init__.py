"""Model export layer: fitted stack (live sklearn object or statically decoded pickle) -> ModelArrays -> flat blob."""
from .blob import BlobError, read_blob, write_blob
from .pickle_static import StaticDecodeError, decode_pickle_bytes, decode_pickle_file
from .walker import (GbtMember, LinearMember, ModelArrays, ModelShapeError, SvcMember, TreeTable,
                     arrays_from_published_pickle, arrays_from_stack, tree_table_from_state)

__all__ = ["BlobError", "read_blob", "write_blob", "StaticDecodeError", "decode_pickle_bytes", "decode_pickle_file",
           "GbtMember", "LinearMember", "ModelArrays", "ModelShapeError", "SvcMember", "TreeTable",
           "arrays_from_published_pickle", "arrays_from_stack", "tree_table_from_state"]
