"""Module interface classes + their registries under the reference's import path
(``deepspeed/inference/v2/modules/interfaces/{attention,embedding,linear,moe,post_norm,pre_norm,unembed}_base.py``)."""
from .implementations import (DSAttentionRegistry, DSEmbeddingBase, DSEmbeddingRegistry, DSLinearBase, DSLinearRegistry,  # noqa: F401
                              DSMoEBase, DSMoERegistry, DSPostNormBase, DSPostNormRegistry, DSPreNormBase, DSPreNormRegistry,
                              DSSelfAttentionBase, DSUnembedBase, DSUnembedRegistry)

DSSelfAttentionRegistry = DSAttentionRegistry


# the reference has this as a package of small files; `__path__` lets its sub-module paths (served from the alias table in
# _import_paths.py) import through this module
__path__: list = []
