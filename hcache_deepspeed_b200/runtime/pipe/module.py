"""PipelineModule: a model expressed as a sequence of layers, partitioned over pipeline stages.

``LayerSpec`` delays construction so each rank only builds the layers of its own stage; ``TiedLayerSpec`` shares
one module (e.g. embeddings) between stages, with its gradients all-reduced over the owning stages.
Partitioning: ``uniform`` (layer count), ``parameters`` (balanced by trainable parameter count), ``type:regex``
(balance layers whose class name matches), ``profile`` is not implemented.
Reference: ``deepspeed/runtime/pipe/module.py`` (LayerSpec :31, TiedLayerSpec :77, PipelineModule :86,
_partition_layers :391, tied comms :455-510)."""
from __future__ import annotations

import os
import re
from functools import partial
from typing import Callable, Iterable, List, Optional

import torch
import torch.nn as nn

from ... import comm as dist
from ...utils.logging import logger
from ..utils import partition_balanced, partition_uniform
from .topology import PipeDataParallelTopology, PipelineParallelGrid, ProcessTopology


class PipelineError(Exception):
    pass


class LayerSpec:
    """Description of a layer: class + constructor arguments; ``build()`` instantiates it."""

    def __init__(self, typename, *module_args, **module_kwargs):
        if not issubclass(typename, nn.Module):
            raise RuntimeError("LayerSpec only supports torch.nn.Module types.")
        self.typename, self.module_args, self.module_kwargs = typename, module_args, module_kwargs
        self.global_rank = dist.get_rank() if dist.is_initialized() else -1

    def __repr__(self):
        from ..utils import call_to_str
        return call_to_str(self.typename.__name__, *self.module_args, **self.module_kwargs)

    def build(self, log: bool = False):
        if log:
            logger.info(f"RANK={self.global_rank} building {self!r}")
        return self.typename(*self.module_args, **self.module_kwargs)


class TiedLayerSpec(LayerSpec):
    def __init__(self, key, typename, *module_args, forward_fn=None, tied_weight_attr=("weight", ), **module_kwargs):
        super().__init__(typename, *module_args, **module_kwargs)
        self.key = key
        self.forward_fn = forward_fn
        self.tied_weight_attr = [tied_weight_attr] if isinstance(tied_weight_attr, str) else list(tied_weight_attr)


class PipelineModule(nn.Module):

    def __init__(self, layers: Iterable, num_stages: Optional[int] = None, topology: Optional[ProcessTopology] = None,
                 loss_fn: Optional[Callable] = None, seed_layers: bool = False, seed_fn=None, base_seed: int = 1234,
                 partition_method: str = "parameters", activation_checkpoint_interval: int = 0,
                 activation_checkpoint_func=None, checkpointable_layers=None, dynamic_shape: bool = False):
        super().__init__()
        if num_stages is None and topology is None:
            raise RuntimeError("must provide num_stages or topology")
        self.micro_offset = 0
        self.loss_fn = loss_fn
        self.checkpointable_layers = checkpointable_layers
        self.seed_layers, self.seed_fn, self.base_seed = seed_layers, seed_fn, base_seed
        self.dynamic_shape = dynamic_shape
        self.world_group = None
        if dist.is_initialized():
            self.global_rank, self.world_size = dist.get_rank(), dist.get_world_size()
        else:
            self.global_rank, self.world_size = 0, 1
        if topology is not None:
            self._topo = topology
            self.num_stages = self._topo.get_dim("pipe")
        else:
            self.num_stages = num_stages
            if self.world_size % num_stages != 0:
                raise RuntimeError(f"num_stages ({num_stages}) must divide distributed world size ({self.world_size})")
            self._topo = PipeDataParallelTopology(num_pp=num_stages, num_dp=self.world_size // num_stages)
        if dist.is_initialized():
            self._grid = PipelineParallelGrid(topology=self._topo)
            self.stage_id = self._grid.get_stage_id()
        else:
            self._grid = None
            self.stage_id = 0
        self._layer_specs = list(layers)
        self._num_layers = len(self._layer_specs)
        self._local_start, self._local_stop = None, None
        self._partition_layers(method=partition_method)
        self.forward_funcs: List[Callable] = []
        self.fwd_map = {}
        self.tied_modules = nn.ModuleDict()
        self.tied_weight_attrs = {}
        self._build()
        self.tied_comms = self._index_tied_modules()
        self._synchronize_tied_weights()
        self.activation_checkpoint_interval = activation_checkpoint_interval
        if activation_checkpoint_func is None:
            from ..activation_checkpointing import checkpointing
            activation_checkpoint_func = checkpointing.checkpoint
        self.activation_checkpoint_func = activation_checkpoint_func

    # ------------------------------------------------------------------------------------------------
    def _count_layer_params(self) -> List[int]:
        counts = [0] * len(self._layer_specs)
        for i, layer in enumerate(self._layer_specs):
            if isinstance(layer, LayerSpec):
                l = layer.build()
                counts[i] = sum(p.numel() for p in l.parameters() if p.requires_grad)
            elif isinstance(layer, nn.Module):
                counts[i] = sum(p.numel() for p in layer.parameters() if p.requires_grad)
        return counts

    def _find_layer_type(self, layername: str) -> List[int]:
        rx = re.compile(layername, re.IGNORECASE)
        idxs = []
        for i, layer in enumerate(self._layer_specs):
            name = layer.typename.__name__ if isinstance(layer, LayerSpec) else \
                layer.__class__.__name__ if isinstance(layer, nn.Module) else getattr(layer, "__name__", "")
            if rx.search(name):
                idxs.append(i)
        if not idxs:
            raise RuntimeError(f"Partitioning '{layername}' found no valid layers to partition.")
        return idxs

    def _partition_layers(self, method: str = "uniform"):
        n = self.num_stages
        method = method.lower()
        if method == "uniform":
            self.parts = partition_uniform(self._num_layers, n)
        elif method == "parameters":
            self.parts = partition_balanced(self._count_layer_params(), n)
        elif method.startswith("type:"):
            w = [0] * self._num_layers
            for i in self._find_layer_type(method.split(":", 1)[1]):
                w[i] = 1
            self.parts = partition_balanced(w, n)
        elif method == "profile":
            raise NotImplementedError(f"Partitioning method {method} not implemented.")
        else:
            raise NotImplementedError(f"Partitioning method {method} not implemented.")
        self._set_bounds(self.parts[self.stage_id], self.parts[self.stage_id + 1])

    def _set_bounds(self, start: int, stop: int):
        self._local_start, self._local_stop = start, stop

    def _build(self):
        for local_idx, layer in enumerate(self._layer_specs[self._local_start:self._local_stop]):
            idx = local_idx + self._local_start
            if self.seed_layers:
                (self.seed_fn or torch.manual_seed)(self.base_seed + idx)
            if isinstance(layer, PipelineModule):
                raise NotImplementedError("RECURSIVE BUILD NOT YET IMPLEMENTED")
            if isinstance(layer, nn.Module):
                name = str(idx)
                self.forward_funcs.append(layer)
                self.fwd_map[name] = len(self.forward_funcs) - 1
                self.add_module(name, layer)
            elif isinstance(layer, TiedLayerSpec):
                if layer.key not in self.tied_modules:
                    self.tied_modules[layer.key] = layer.build()
                    self.tied_weight_attrs[layer.key] = layer.tied_weight_attr
                if layer.forward_fn is None:
                    self.forward_funcs.append(self.tied_modules[layer.key])
                else:
                    self.forward_funcs.append(partial(layer.forward_fn, self.tied_modules[layer.key]))
            elif isinstance(layer, LayerSpec):
                module = layer.build()
                name = str(idx)
                self.forward_funcs.append(module)
                self.fwd_map[name] = len(self.forward_funcs) - 1
                self.add_module(name, module)
            else:
                self.forward_funcs.append(layer)  # plain callable
        for p in self.parameters():
            p.ds_pipe_replicated = False

    # ------------------------------------------------------------------------------------------------
    def _is_checkpointable(self, funcs) -> bool:
        if self.checkpointable_layers is not None:
            return all(f.__class__.__name__ in self.checkpointable_layers for f in funcs)
        params = [f.parameters() for f in funcs if isinstance(f, nn.Module)]
        return any(len(list(p)) > 0 for p in params)

    def forward(self, forward_input):
        self.micro_offset += 1

        def exec_range(start, end):
            def run(*inputs):
                x = inputs[0] if len(inputs) == 1 else inputs
                for f in self.forward_funcs[start:end]:
                    x = f(x)
                return x
            return run

        if self.activation_checkpoint_interval == 0:
            return exec_range(0, len(self.forward_funcs))(forward_input)
        x = forward_input
        n = len(self.forward_funcs)
        for start in range(0, n, self.activation_checkpoint_interval):
            end = min(start + self.activation_checkpoint_interval, n)
            funcs = self.forward_funcs[start:end]
            if not isinstance(x, tuple):
                x = (x, )
            # a segment whose inputs carry no gradient (token ids into the embedding of the first stage) is run plainly: a
            # re-entrant checkpoint there would detach its parameters from the graph and they would never get a gradient
            feeds_grad = any(torch.is_tensor(t) and t.requires_grad for t in x)
            if self._is_checkpointable(funcs) and (feeds_grad or not torch.is_grad_enabled()):
                x = self.activation_checkpoint_func(exec_range(start, end), *x)
            else:
                x = exec_range(start, end)(*x)
        return x

    # ---- tied layers -----------------------------------------------------------------------------------
    def _index_tied_modules(self):
        tied = {}
        if self._grid is None or self._topo.get_dim("pipe") == 1:
            return tied
        specs = self._layer_specs
        keys = sorted({s.key for s in specs if isinstance(s, TiedLayerSpec)})
        for key in keys:
            layer_idxs = [i for i, s in enumerate(specs) if isinstance(s, TiedLayerSpec) and s.key == key]
            stages = sorted({self.stage_owner(i) for i in layer_idxs})
            for dp in range(self._grid.data_parallel_size):
                for mp in range(self._grid.get_slice_parallel_world_size()):
                    ranks = []
                    for s in stages:
                        if self._grid.get_slice_parallel_world_size() > 1:
                            ranks.append(self._grid.stage_to_global(stage_id=s, data=dp, model=mp))
                        else:
                            ranks.append(self._grid.stage_to_global(stage_id=s, data=dp))
                    group = dist.new_group(ranks=ranks)
                    if self.global_rank in ranks:
                        assert key in self.tied_modules
                        tied[key] = {"ranks": ranks, "group": group, "weight_attr": self.tied_weight_attrs[key],
                                     "module": self.tied_modules[key]}
                        if self.global_rank != ranks[0]:
                            for p in self.tied_modules[key].parameters():
                                p.ds_pipe_replicated = True
        return tied

    def _tied_weight(self, module, attr):
        obj = module
        for part in attr.split("."):
            obj = getattr(obj, part)
        return obj

    def _synchronize_tied_weights(self):
        for key, comm in self.tied_comms.items():
            for attr in comm["weight_attr"]:
                dist.broadcast(self._tied_weight(comm["module"], attr).data, src=min(comm["ranks"]), group=comm["group"])

    def allreduce_tied_weight_gradients(self):
        for key, comm in self.tied_comms.items():
            for attr in comm["weight_attr"]:
                w = self._tied_weight(comm["module"], attr)
                if not w.requires_grad:
                    continue
                if w.grad is None:      # every stage that holds the tied weight must join the collective
                    w.grad = torch.zeros_like(w)
                dist.all_reduce(w.grad, group=comm["group"])

    # ---- checkpoints: one file per layer, independent of how the layers are cut into stages -------------------------------
    # layer_<II>-model_<MM>-model_states.pt (ref: pipe/module.py ckpt_layer_path / save_state_dict / load_state_dir)
    def _layer_module(self, global_idx: int):
        """The module that holds layer ``global_idx``'s parameters on this stage (a tied layer resolves to its shared module)."""
        spec = self._layer_specs[global_idx]
        if isinstance(spec, TiedLayerSpec):
            return self.tied_modules[spec.key] if spec.key in self.tied_modules else None
        f = self.forward_funcs[global_idx - self._local_start]
        return f if isinstance(f, nn.Module) else None

    def ckpt_layer_path(self, ckpt_dir: str, global_idx: int) -> str:
        mp = self._grid.get_slice_parallel_rank() if self._grid is not None else 0
        return os.path.join(ckpt_dir, f"layer_{global_idx:02d}-model_{mp:02d}-model_states.pt")

    def ckpt_layer_path_list(self, ckpt_dir: str, local_layer_idx: int) -> List[str]:
        """Every model-parallel slice file of this stage's ``local_layer_idx``-th layer, sorted."""
        import glob
        idx = local_layer_idx + self._local_start
        return sorted(glob.glob(os.path.join(ckpt_dir, f"layer_{idx:02d}-*model_states.pt")))

    def set_checkpoint_interval(self, interval: int) -> None:
        """Layers per activation-checkpointed segment (0 = no activation checkpointing)."""
        assert interval >= 0
        self.activation_checkpoint_interval = interval

    @property
    def checkpoint_interval(self) -> int:
        return self.activation_checkpoint_interval

    def get_additional_losses(self):
        """Override to report named side losses (``{"name": tensor}``) next to the main loss; ``None`` = none."""
        return None

    def compile(self, *args, **kwargs) -> None:
        """The reference wraps each layer in ``torch.compile`` here (pipe/module.py:687). This framework has no tracing
        compiler in its hot path: the stage's layers run as written (their fused CUDA ops are already single launches)
        and the call is accepted so scripts that issue it keep working."""
        self._compile_requested = (args, kwargs)

    def save_state_dict(self, save_dir: str, checkpoint_engine, exclude_frozen_params: bool = False) -> None:
        dp_rank = self._grid.data_parallel_id if self._grid is not None else 0
        if dp_rank != 0:
            return
        os.makedirs(save_dir, exist_ok=True)
        for idx in range(self._local_start, self._local_stop):
            m = self._layer_module(idx)
            if m is None or not any(True for _ in m.state_dict()):
                continue
            sd = {k: v.detach().cpu().clone() for k, v in m.state_dict().items()}
            if exclude_frozen_params:
                frozen = {n for n, p in m.named_parameters() if not p.requires_grad}
                sd = {k: v for k, v in sd.items() if k not in frozen}
            checkpoint_engine.save(sd, self.ckpt_layer_path(save_dir, idx))

    def load_state_dir(self, load_dir: str, checkpoint_engine, strict: bool = True) -> None:
        for idx in range(self._local_start, self._local_stop):
            m = self._layer_module(idx)
            if m is None or not any(True for _ in m.state_dict()):
                continue
            path = self.ckpt_layer_path(load_dir, idx)
            if not os.path.exists(path):
                if strict:
                    raise FileNotFoundError(f"pipeline checkpoint file {path} is missing")
                continue
            m.load_state_dict(checkpoint_engine.load(path, map_location="cpu"), strict=strict)
        self._synchronize_tied_weights()

    def get_tied_weights_and_groups(self):
        return [(self._tied_weight(c["module"], a), c["group"]) for c in self.tied_comms.values() for a in c["weight_attr"]]

    # ---- queries -----------------------------------------------------------------------------------------
    def stage_owner(self, layer_idx: int) -> int:
        assert 0 <= layer_idx < self._num_layers
        for stage in range(self._topo.get_dim("pipe")):
            if self.parts[stage] <= layer_idx < self.parts[stage + 1]:
                return stage
        raise RuntimeError(f"Layer {layer_idx} not owned? parts={self.parts}")

    def partitions(self):
        return self.parts

    def topology(self):
        return self._topo

    def mpu(self):
        return self._grid

    def num_pipeline_stages(self):
        return self._topo.get_dim("pipe")

    def ckpt_prefix(self, checkpoints_path, tag):
        import os
        return os.path.join(checkpoints_path, str(tag), "module")

