"""AutoTP: tensor parallelism for a model WITHOUT a hand-written policy.

The parser walks each transformer block, finds the Linear layers whose output feeds the residual stream (the last
linear of the attention and of the MLP: ``o_proj``, ``out_proj``, ``dense``, ``down_proj``, ``fc2``, ``c_proj``, ...)
— those become row-parallel ``LinearAllreduce`` layers; every other Linear in the block becomes column-parallel
``LinearLayer``. Head-count attributes of the attention modules are divided by the TP degree. Fused QKV and fused
gate/up weights are sliced consistently (``fusedqkv_utils``), the LM head becomes ``LmHeadLinearAllreduce``.
Reference: ``deepspeed/module_inject/auto_tp.py`` (ReplaceWithTensorSlicing :32, Loading :133, AutoTP :193,
tp_parser :285, _replace :348, update_mp_params :447).
"""
from __future__ import annotations

from typing import Dict, List, Optional, Tuple

import torch
from torch import nn

from .. import comm as dist
from ..utils.logging import logger
from .fusedqkv_utils import fused_qkv_layout, prepare_tp_fused_qkvw, require_tp_fused_qkvw, shard_chunk_mlp
from .layers import LinearAllreduce, LinearLayer, LmHeadLinearAllreduce
from .tp_shard import get_shard_size, get_shard_size_list, set_n_embd, set_num_kv_heads

ALL_REDUCE_NAMES = ("o_proj", "out_proj", "dense_4h_to_h", "down_proj", "fc2", "fc_out", "c_proj", "attention.dense",
                    "self_attention.dense", "self_attn.dense", "wo", "w2", "dense_out", "attn.proj", "mlp.dense_4h_to_h",
                    "attention.output.dense", "output.dense")
UNSUPPORTED = ("deberta", "flaubert", "fsmt", "gpt2-as-conv1d", "led", "longformer", "xlm", "xlnet")
HEAD_ATTRS = ("num_heads", "num_attention_heads", "num_attn_heads", "n_heads", "n_head", "num_kv_heads", "num_key_value_heads",
              "num_kv", "all_head_size", "embed_dim", "hidden_size", "n_embd", "split_size", "num_key_value_groups")


def is_linear(m: nn.Module) -> bool:
    """nn.Linear, or HuggingFace's ``Conv1D`` (GPT-2: the same affine map with the weight stored [in, out])."""
    return isinstance(m, nn.Linear) or (type(m).__name__ == "Conv1D" and hasattr(m, "nf") and hasattr(m, "weight"))


def linear_weight_bias(m: nn.Module):
    """(weight [out, in], bias) of a linear-like module."""
    w = m.weight.data
    if not isinstance(m, nn.Linear):
        w = w.t()
    return w, (m.bias.data if getattr(m, "bias", None) is not None else None)


class ReplaceWithTensorSlicing:
    """Copies the rank's slice of a source tensor into a (smaller) destination — row, column or fused-QKV split."""

    def __init__(self, mp_group=None, mp_size: int = 1, out_dim: int = 1, in_dim: int = 0):
        self.gpu_index = dist.get_rank(group=mp_group) if (mp_group is not None and dist.is_initialized()) else 0
        self.mp_size, self.out_dim, self.in_dim = mp_size, out_dim, in_dim

    def merge_assert(self, dim1, dim2):
        assert dim1 > dim2, "Merging tensors is not allowed here! Please use deepspeed load_checkpoint for merging your checkpoints"

    def strided_copy(self, dst: torch.Tensor, src: torch.Tensor, num_splits: int, int8: bool = False,
                     allocate_tensor: bool = False) -> torch.Tensor:
        """Fused tensors made of ``num_splits`` equal parts (q|k|v): take this rank's slice of every part."""
        if src is None:
            return src
        if src.shape == dst.shape:
            return src if not allocate_tensor else src.clone()
        outer = src.dim() - 1 if src.dim() > 1 and src.shape[-1] != dst.shape[-1] else 0
        parts = torch.chunk(src, num_splits, dim=outer)
        sl = []
        for p in parts:
            sizes = get_shard_size_list(p.shape[outer], self.mp_size)
            start = sum(sizes[:self.gpu_index])
            sl.append(p.narrow(outer, start, sizes[self.gpu_index]))
        out = torch.cat(sl, dim=outer)
        dst.data.copy_(out.reshape(dst.shape))
        return dst

    def copy(self, dst: torch.Tensor, src: torch.Tensor, int8: bool = False, allocate_tensor: bool = False) -> torch.Tensor:
        if src is None:
            return src
        if src.shape == dst.shape:
            dst.data.copy_(src)
            return dst
        for dim in range(src.dim()):
            if src.shape[dim] != dst.shape[dim]:
                self.merge_assert(src.shape[dim], dst.shape[dim])
                sizes = get_shard_size_list(src.shape[dim], self.mp_size)
                start = sum(sizes[:self.gpu_index])
                dst.data.copy_(src.narrow(dim, start, sizes[self.gpu_index]))
                return dst
        return dst


class Loading:

    @staticmethod
    def is_load_module(module) -> bool:
        load_layers = (nn.Linear, nn.Embedding, nn.LayerNorm)
        names = ("LlamaRMSNorm", "RMSNorm", "MistralRMSNorm", "T5LayerNorm", "OPTLearnedPositionalEmbedding", "Phi3RMSNorm",
                 "Qwen2RMSNorm")
        return isinstance(module, load_layers) or type(module).__name__ in names

    @staticmethod
    def load_buffer(module, state_dict, prefix):
        for name in module._buffers.keys():
            key = prefix + name
            if key in state_dict and module._buffers[name] is not None:
                module._buffers[name].data.copy_(state_dict[key])

    @staticmethod
    def load(module, state_dict, prefix, mp_group=None):
        slicer = ReplaceWithTensorSlicing(mp_group=mp_group, mp_size=dist.get_world_size(group=mp_group) if mp_group is not None else 1)
        for pname in ("weight", "bias"):
            key = prefix + pname
            p = getattr(module, pname, None)
            if key in state_dict and p is not None:
                slicer.copy(p.data, state_dict[key])


class AutoTP:

    def __init__(self, module, all_reduce_linears, prefix, state_dict, linear_layer_setting, orig_layer_impl,
                 keep_module_on_host: bool = False):
        self.module, self.all_reduce_linears, self.prefix = module, all_reduce_linears, prefix
        self.state_dict, self.linear_layer_setting, self.orig_layer_impl = state_dict, linear_layer_setting, orig_layer_impl
        self.mp_size, self.mp_group, self.globals_rank = None, None, 0
        self.keep_module_on_host = keep_module_on_host
        self.conv_linear_layer = False
        self.default_heads = (None, None)       # (query heads, KV heads) of the model config: for attention modules that
                                                # keep no head-count attribute of their own (GPT-NeoX)

    # ---- discovery -------------------------------------------------------------------------------------
    @staticmethod
    def in_module_list(module, module_list) -> bool:
        return any(type(item).__name__ == type(module).__name__ for item in module_list)

    @staticmethod
    def get_module_list(model) -> List[nn.Module]:
        """One representative of every repeated block type (the children of ModuleLists)."""
        out = []
        for child in model.children():
            if isinstance(child, nn.ModuleList):
                for m in child.children():
                    if not out or not AutoTP.in_module_list(m, out):
                        out.append(m)
            else:
                out += [m for m in AutoTP.get_module_list(child) if not AutoTP.in_module_list(m, out)]
        return out

    @staticmethod
    def kernel_supported(module_list) -> bool:
        """True when one of the block types has a kernel-injection policy (then injection is the better route)."""
        from .replace_policy import policy_class_names
        return any(type(m).__name__ in policy_class_names for m in module_list)     # block class name -> family

    def get_model_num_kv_heads(self, config):
        """KV-head count under the names model configs use for it (falls back to the query-head count)."""
        for name in ("multi_query_group_num", "num_kv_heads", "num_key_value_heads", "num_attention_heads", "n_heads",
                     "attention_heads"):
            v = getattr(config, name, None)
            if v is not None:
                return v
        return None

    def update_linear_policies(self) -> None:
        """Which layer types ``_replace_module`` shards. ``linear_layer_setting`` (one or two types: linear[, embedding])
        overrides the default of ``nn.Linear`` (+ HuggingFace ``Conv1D`` for GPT-2 blocks)."""
        self.conv_linear_layer = False
        if self.linear_layer_setting is not None:
            setting = self.linear_layer_setting if isinstance(self.linear_layer_setting, (list, tuple)) else [self.linear_layer_setting]
            self.linear_policies = {setting[0]: self._replace}
            if len(setting) == 2:
                self.linear_policies[setting[1]] = None      # embeddings stay replicated here (vocab-parallel is opt-in)
        else:
            self.linear_policies = {nn.Linear: self._replace}
            if self.orig_layer_impl is not None and type(self.orig_layer_impl).__name__ == "type" and \
                    self.orig_layer_impl.__name__ == "GPT2Block":
                self.conv_linear_layer = True
            try:
                from transformers.pytorch_utils import Conv1D
                self.linear_policies[Conv1D] = self._replace
            except ImportError:
                pass

    @staticmethod
    def supported(model) -> bool:
        name = str(type(model)).lower()
        return not any(u in name for u in UNSUPPORTED)

    @staticmethod
    def get_layers(parent, module) -> List[str]:
        return [parent + "." + n if parent else n for n, m in module.named_modules() if isinstance(m, nn.Linear) and n]

    @staticmethod
    def update_policy_list(policy_list, new_module, new_gems):
        for i, (mod, gems) in enumerate(policy_list):
            if type(mod).__name__ == type(new_module).__name__:
                policy_list[i] = (mod, list(dict.fromkeys(list(gems) + list(new_gems))))
                return policy_list
        policy_list.append((new_module, list(new_gems)))
        return policy_list

    @staticmethod
    def tp_parser(model) -> List[Tuple[nn.Module, List[str]]]:
        """[(block, [names of its linears that need an all-reduce])]. A linear is row-parallel when its name says so
        (``o_proj``, ``fc2``, ``c_proj``, ...); in a sub-module with several linears and no such name the LAST one is taken
        (the projection back to the residual stream: Phi's ``self_attn.dense``, GPT-J's ``mlp.fc_out`` style layouts)."""
        assert AutoTP.supported(model), "AutoTP not supported for model. Please use kernel injection since container policy for model exists."
        policy_list = []
        for block in AutoTP.get_module_list(model):
            linears = [n for n, m in block.named_modules() if is_linear(m) and n]
            gems = [n for n in linears if any(n.endswith(k) or n.split(".")[-1] == k for k in ALL_REDUCE_NAMES)]
            for cname, child in block.named_children():
                cl = [n for n, m in child.named_modules() if is_linear(m) and n]
                if len(cl) >= 2 and not any(g.startswith(cname + ".") for g in gems):
                    gems.append(cname + "." + cl[-1])
            if gems:
                AutoTP.update_policy_list(policy_list, block, gems)
        assert policy_list, "AutoTP not supported for model: no row-parallel linears found. Please use kernel injection."
        return policy_list

    # ---- replacement -----------------------------------------------------------------------------------
    def set_tensor_parallel_config(self, mp_size, mp_group):
        self.mp_size, self.mp_group = mp_size, mp_group
        self.rank = dist.get_rank(group=mp_group) if (mp_group is not None and dist.is_initialized()) else 0

    def _slice(self, t: Optional[torch.Tensor], dim: int, name: str):
        if t is None:
            return None
        sizes = get_shard_size_list(t.shape[dim], self.mp_size, name)
        start = sum(sizes[:self.rank])
        return t.narrow(dim, start, sizes[self.rank]).clone().contiguous()

    def _replace(self, child: nn.Module, name: str, full_name: str, block) -> nn.Module:
        if getattr(child, "replaced", False):
            return child
        w, b = linear_weight_bias(child)
        if any(full_name.endswith(g) or name == g for g in self.all_reduce_linears):
            new = LinearAllreduce(self._slice(w, 1, full_name), b.clone() if b is not None else None, self.mp_group)
        elif require_tp_fused_qkvw(full_name, self.mp_size):
            heads = self._heads_of(block)
            kind = fused_qkv_layout(block)
            nw = prepare_tp_fused_qkvw(kind, w, self.mp_size, self.rank, heads[0], heads[1]).clone()
            nb = prepare_tp_fused_qkvw(kind, b, self.mp_size, self.rank, heads[0], heads[1]).clone() if b is not None else None
            new = LinearLayer(weight=nw, bias=nb, mp_group=self.mp_group)
        elif any(k in full_name for k in ("gate_up_proj", "dense_h_to_4h_fused")):
            nw, nb = shard_chunk_mlp(w, b, self.rank, self.mp_size)
            new = LinearLayer(weight=nw.clone(), bias=None if nb is None else nb.clone(), mp_group=self.mp_group)
        else:
            new = LinearLayer(weight=self._slice(w, 0, full_name), bias=self._slice(b, 0, full_name), mp_group=self.mp_group)
        new.replaced = True
        return new

    def _heads_of(self, block):
        """(query heads, KV heads) of the block BEFORE sharding (``update_mp_params`` keeps the originals)."""
        cfg_heads, cfg_kv = None, None
        for m in block.modules():
            orig = getattr(m, "_autotp_orig", None)
            if orig:
                for a in ("num_heads", "num_attention_heads", "n_head", "n_heads"):
                    if a in orig and cfg_heads is None:
                        cfg_heads = orig[a]
                for a in ("num_key_value_heads", "num_kv_heads", "num_kv"):
                    if a in orig and cfg_kv is None:
                        cfg_kv = orig[a]
        if cfg_heads is not None:
            return cfg_heads, cfg_kv
        for m in block.modules():
            for a in ("num_heads", "num_attention_heads", "n_head", "n_heads"):
                if isinstance(getattr(m, a, None), int) and cfg_heads is None:
                    cfg_heads = getattr(m, a)
            for a in ("num_key_value_heads", "num_kv_heads", "num_kv"):
                if isinstance(getattr(m, a, None), int) and cfg_kv is None:
                    cfg_kv = getattr(m, a)
        if cfg_heads is None:
            cfg_heads = self.default_heads[0]
        if cfg_kv is None:
            cfg_kv = self.default_heads[1]
        return cfg_heads, cfg_kv

    def update_mp_params(self, child):
        """Divide head-count / width attributes of attention modules by the TP degree."""
        if getattr(child, "_autotp_updated", False):
            return
        child._autotp_orig = {a: getattr(child, a) for a in HEAD_ATTRS if isinstance(getattr(child, a, None), int)}
        for a in HEAD_ATTRS:
            v = getattr(child, a, None)
            if isinstance(v, int) and a != "num_key_value_groups":
                if a in ("num_kv_heads", "num_key_value_heads", "num_kv") and v < self.mp_size:
                    continue        # fewer KV heads than ranks (multi-query): K / V stay replicated
                try:
                    setattr(child, a, get_shard_size(v, self.mp_size, rank=self.rank))
                except AttributeError:
                    pass
        # modules that read head counts from their (shared) config at run time (Phi-3: ``config.num_attention_heads`` in
        # forward) get a private shallow copy of it with this rank's counts
        cfg = getattr(child, "config", None)
        if cfg is not None and any(isinstance(getattr(cfg, a, None), int) for a in ("num_attention_heads", "n_head")):
            import copy
            local = copy.copy(cfg)
            for a in ("num_attention_heads", "n_head", "num_key_value_heads", "num_kv_heads"):
                v = getattr(cfg, a, None)
                if isinstance(v, int) and not (a in ("num_key_value_heads", "num_kv_heads") and v < self.mp_size):
                    try:
                        setattr(local, a, get_shard_size(v, self.mp_size, rank=self.rank))
                    except AttributeError:
                        pass
            try:
                child.config = local
            except AttributeError:
                pass
        child._autotp_updated = True

    def _replace_module(self, r_module, block=None, prev_name: str = "", prev_class_name: str = ""):
        for name, child in list(r_module.named_children()):
            full = f"{prev_name}.{name}" if prev_name else name
            if is_linear(child):
                setattr(r_module, name, self._replace(child, name, full, block if block is not None else r_module))
            elif isinstance(child, nn.Embedding) and False:
                pass
            else:
                has_counts = any(hasattr(child, a) for a in ("num_heads", "num_attention_heads", "n_head", "n_heads"))
                if not has_counts and "attention" in type(child).__name__.lower().replace("attn", "attention"):
                    # attention modules that keep only KV-head counts / read the rest from their config (Phi-3)
                    has_counts = any(hasattr(child, a) for a in ("num_key_value_heads", "num_kv_heads")) or \
                        isinstance(getattr(getattr(child, "config", None), "num_attention_heads", None), int)
                if has_counts and any(is_linear(m) for m in child.children()):
                    self.update_mp_params(child)
                self._replace_module(child, block if block is not None else r_module, full, type(child).__name__)
        return r_module

    def _replace_last_linear_module(self, model):
        for name in ("lm_head", "embed_out"):
            head = getattr(model, name, None)
            if isinstance(head, nn.Linear) and not getattr(head, "replaced", False):
                w = head.weight.data
                sizes = get_shard_size_list(w.shape[1], self.mp_size, name)
                start = sum(sizes[:self.rank])
                new = LmHeadLinearAllreduce(w[:, start:start + sizes[self.rank]].clone(), self.rank, self.mp_size,
                                            head.bias.data.clone() if head.bias is not None else None, self.mp_group)
                new.replaced = True
                setattr(model, name, new)
        return model

    # ---- entry point -----------------------------------------------------------------------------------
    @staticmethod
    def apply(model: nn.Module, mp_size: int, rank: int, mp_group, replace_lm_head: bool = True,
              injection_policy: Optional[Dict] = None) -> nn.Module:
        """Shard ``model`` in place over ``mp_group`` (``mp_size`` ranks, this process is ``rank``).
        ``injection_policy``: ``{BlockClass: ("attn.out_proj", "mlp.fc2")}`` — the user names the row-parallel linears of
        each block class instead of the parser guessing them (``init_inference(..., injection_policy=...)``,
        ref inference/engine.py:147-165)."""
        if mp_size <= 1:
            return model
        cfg = getattr(model, "config", None)
        kv = None
        for a in ("num_key_value_heads", "num_kv_heads", "num_attention_heads", "n_head", "n_heads"):
            if cfg is not None and isinstance(getattr(cfg, a, None), int):
                kv = getattr(cfg, a)
                break
        if kv is not None and 1 < kv < mp_size:
            # every rank raises the same error before any collective: a grouped-query model cannot give each rank a whole KV
            # head here (multi-query models, kv = 1, keep K / V replicated instead)
            raise ValueError(f"AutoTP: {kv} key/value heads cannot be split over {mp_size} ranks; use tp_size <= {kv}, or kernel "
                             f"injection (the ragged executor replicates KV heads across ranks)")
        set_num_kv_heads(kv)
        if cfg is not None and isinstance(getattr(cfg, "hidden_size", None), int):
            set_n_embd(cfg.hidden_size)
        n_q = next((getattr(cfg, a) for a in ("num_attention_heads", "n_head", "num_heads", "n_heads")
                    if cfg is not None and isinstance(getattr(cfg, a, None), int)), None)
        if injection_policy:
            by_type: Dict[str, List[str]] = {}
            for cls, names in injection_policy.items():
                names = [names] if isinstance(names, str) else list(names)
                by_type[cls if isinstance(cls, str) else cls.__name__] = names
            present = {type(m).__name__ for m in model.modules()}
            missing = [k for k in by_type if k not in present]
            if missing:
                raise ValueError(f"injection_policy names module classes the model does not contain: {missing}")
        else:
            policies = AutoTP.tp_parser(model)
            by_type = {type(m).__name__: gems for m, gems in policies}
        for module in model.modules():
            tname = type(module).__name__
            if tname in by_type and not getattr(module, "_autotp_done", False):
                at = AutoTP(model, by_type[tname], "", None, None, type(module))
                at.set_tensor_parallel_config(mp_size, mp_group)
                at.rank = rank
                at.default_heads = (n_q, kv)
                at._replace_module(module)
                module._autotp_done = True
        _patch_alibi_builders(model, mp_size, rank)
        if replace_lm_head:
            at = AutoTP(model, [], "", None, None, None)
            at.set_tensor_parallel_config(mp_size, mp_group)
            at.rank = rank
            at._replace_last_linear_module(model)
        logger.info(f"AutoTP: sharded {sum(len(g) for g in by_type.values())} row-parallel linear kinds over {mp_size} ranks")
        return model


def _patch_alibi_builders(model: nn.Module, mp_size: int, rank: int) -> None:
    """BLOOM-style models build an ALiBi tensor for ALL heads in the model's forward; after sharding every rank needs the
    rows of its own heads (ref: auto_tp_model_utils.py build_bloom_alibi_tensor, replace_module.py:309-313)."""
    from .auto_tp_model_utils import build_bloom_alibi_tensor
    for m in model.modules():
        if hasattr(m, "build_alibi_tensor") and type(m).__name__ in ("BloomModel", ):
            def build(attention_mask, num_heads, dtype, _m=m):
                return build_bloom_alibi_tensor(attention_mask, num_heads, dtype, mp_size=mp_size, mp_rank=rank)
            m.build_alibi_tensor = build


def move(tensor: torch.Tensor, device, copy: bool = True) -> torch.Tensor:
    """A shard on its target device as its OWN storage (``copy=True`` even for cpu -> cpu), so that the full tensor it was
    cut from can be freed; meta tensors become uninitialised tensors of the same shape."""
    if tensor.is_meta:
        return torch.empty_like(tensor, device=device)
    return tensor.to(device, copy=copy)
