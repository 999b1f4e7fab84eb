"""``InferenceV2Policy`` / ``ContainerMap`` (``deepspeed/inference/v2/model_implementations/inference_policy_base.py``)."""
from __future__ import annotations

from typing import Dict, Iterable, List, Optional, Tuple

from . import MODEL_TYPE_TO_POLICY, POLICIES, InferenceV2Policy  # noqa: F401

POLICIES = POLICIES


class ContainerMap:
    """Checkpoint-name prefixes -> layer containers. The reference routes every checkpoint tensor through this map; the
    weight assembler of this framework (``models/weights.py``) does the routing itself from the architecture's name
    table, so the map is a thin, order-preserving registry kept for policies written against the reference."""

    def __init__(self) -> None:
        self._prefix_map: Dict[str, object] = {}
        self._transformer_prefixes: Optional[List[str]] = None
        self._transformer_containers: Optional[Iterable] = None
        self._non_transformer = None
        self._unmapped: List[str] = []

    def set_transformer_params(self, prefixes, containers) -> None:
        self._transformer_prefixes = [prefixes] if isinstance(prefixes, str) else list(prefixes)
        self._transformer_containers = list(containers)

    def set_non_transformer_params(self, container) -> None:
        self._non_transformer = container

    def set_unmapped_params(self, prefixes) -> None:
        self._unmapped = [prefixes] if isinstance(prefixes, str) else list(prefixes)

    @property
    def transformer_params(self):
        return self._transformer_containers

    @property
    def non_transformer_params(self):
        return self._non_transformer

    def resolve(self, name: str) -> Tuple[Optional[object], Optional[str]]:
        """``(container, name inside the container)`` for a checkpoint tensor name; ``(None, None)`` for unmapped ones."""
        if any(name.startswith(u) for u in self._unmapped):
            return None, None
        for prefix in self._transformer_prefixes or []:
            if name.startswith(prefix + "."):
                rest = name[len(prefix) + 1:]
                idx, _, tail = rest.partition(".")
                if idx.isdigit() and self._transformer_containers is not None:
                    return self._transformer_containers[int(idx)], tail
        if self._non_transformer is not None:
            return self._non_transformer, name
        raise ValueError(f"Cannot find container for {name}, please double check the Containers/ContainerMap")

    def map_param(self, name: str, tensor) -> None:
        container, inner = self.resolve(name)
        if container is not None:
            container.set_dependency(inner, tensor)

    def validate(self) -> None:
        for c in list(self._transformer_containers or []) + ([self._non_transformer] if self._non_transformer is not None else []):
            if hasattr(c, "is_populated") and not c.is_populated:
                raise ValueError(f"container {type(c).__name__} was not fully populated: {getattr(c, 'missing', lambda: '')()}")
