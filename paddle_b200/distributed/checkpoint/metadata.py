"""Sharded-checkpoint metadata (reference: python/paddle/distributed/checkpoint/metadata.py:19-43).  The classes
pickle under the reference's module path so ``.metadata`` files interoperate."""
from __future__ import annotations

from dataclasses import dataclass, field


@dataclass
class LocalTensorMetadata:
    global_offset: tuple
    local_shape: tuple
    dtype: str = ""


@dataclass(frozen=True)
class LocalTensorIndex:
    tensor_key: str
    global_offset: tuple


@dataclass
class Metadata:
    state_dict_metadata: dict = field(default_factory=dict)   # key -> [LocalTensorMetadata]
    storage_metadata: dict = field(default_factory=dict)      # LocalTensorIndex -> file name
    flat_mapping: dict = field(default_factory=dict)          # flat key -> tuple path in the nested dict


for _c in (LocalTensorMetadata, LocalTensorIndex, Metadata):
    _c.__module__ = "paddle.distributed.checkpoint.metadata"
