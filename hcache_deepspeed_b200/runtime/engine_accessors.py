"""The engine's configuration getters and small utility methods, as a mixin of ``DeepSpeedEngine``.

The reference exposes its configuration through ~110 one-line getters on the engine (``deepspeed/runtime/engine.py:480-960``)
which user code and integrations (HF Trainer, Megatron-DeepSpeed, DeepSpeed-Chat) call freely. Here they are generated from
one table (name -> path into ``DeepSpeedConfig``), so there is a single place to read what each one returns; the handful
of getters that compute something are written out below the table, followed by the gradient-bucket helpers
(``allreduce_bucket`` & co., ``engine.py:2560-2720``) and dataset predicates.
"""
from __future__ import annotations

import os
from typing import Any, Callable, Dict, Optional, Sequence

import torch

from .. import comm as dist
from .utils import flatten_dense_tensors, unflatten_dense_tensors

_MISSING = object()


def _dig(root: Any, path: Sequence[str], default: Any = None) -> Any:
    """Follow ``path`` through attributes / mapping keys; ``default`` as soon as a hop is absent."""
    cur = root
    for hop in path:
        if cur is None:
            return default
        if isinstance(cur, dict):
            cur = cur.get(hop, _MISSING)
        else:
            cur = getattr(cur, hop, _MISSING)
        if cur is _MISSING:
            return default
    return cur


# getter name -> (path below engine._config, default when absent)
_CONFIG_GETTERS: Dict[str, tuple] = {
    "aio_config": (("aio_config", ), None),
    "amp_params": (("amp_params", ), None),
    "autotp_size": (("tensor_parallel_config", "autotp_size"), 0),
    "tensor_parallel_config": (("tensor_parallel_config", ), None),
    "autotuning_end_profile_step": (("autotuning_config", "end_profile_step"), 5),
    "autotuning_start_profile_step": (("autotuning_config", "start_profile_step"), 3),
    "autotuning_metric": (("autotuning_config", "metric"), "throughput"),
    "checkpoint_tag_validation_enabled": (("checkpoint_tag_validation_enabled", ), True),
    "checkpoint_tag_validation_fail": (("checkpoint_tag_validation_fail", ), False),
    "curriculum_enabled_legacy": (("curriculum_enabled_legacy", ), False),
    "curriculum_params_legacy": (("curriculum_params_legacy", ), False),
    "data_efficiency_config": (("data_efficiency_config", ), None),
    "data_efficiency_enabled": (("data_efficiency_enabled", ), False),
    "data_sampling_config": (("data_efficiency_config", "data_sampling"), None),
    "data_sampling_enabled": (("data_efficiency_config", "data_sampling", "enabled"), False),
    "curriculum_learning_config": (("data_efficiency_config", "data_sampling", "curriculum_learning"), None),
    "curriculum_learning_enabled": (("data_efficiency_config", "data_sampling", "curriculum_learning", "enabled"), False),
    "random_ltd_config": (("data_efficiency_config", "data_routing", "random_ltd"), None),
    "random_ltd_enabled": (("data_efficiency_config", "data_routing", "random_ltd", "enabled"), False),
    "dataloader_drop_last": (("dataloader_drop_last", ), False),
    "dump_state": (("dump_state", ), False),
    "dynamic_loss_scale_args": (("dynamic_loss_scale_args", ), None),
    "eigenvalue_enabled": (("eigenvalue_config", "enabled"), False),
    "eigenvalue_gas_boundary_resolution": (("eigenvalue_config", "gas_boundary_resolution"), 1),
    "eigenvalue_layer_name": (("eigenvalue_config", "layer_name"), "bert.encoder.layer"),
    "eigenvalue_layer_num": (("eigenvalue_config", "layer_num"), 0),
    "eigenvalue_max_iter": (("eigenvalue_config", "max_iter"), 100),
    "eigenvalue_stability": (("eigenvalue_config", "stability"), 1e-6),
    "eigenvalue_tol": (("eigenvalue_config", "tol"), 1e-2),
    "eigenvalue_verbose": (("eigenvalue_config", "verbose"), False),
    "elasticity_enabled": (("elasticity_enabled", ), False),
    "flops_profiler_module_depth": (("flops_profiler_config", "module_depth"), -1),
    "flops_profiler_output_file": (("flops_profiler_config", "output_file"), None),
    "flops_profiler_recompute_fwd_factor": (("flops_profiler_config", "recompute_fwd_factor"), 0.0),
    "flops_profiler_top_modules": (("flops_profiler_config", "top_modules"), 1),
    "fp16_auto_cast": (("fp16_config", "auto_cast"), False),
    "fp16_master_weights_and_gradients": (("fp16_config", "fp16_master_weights_and_grads"), False),
    "graph_harvesting": (("graph_harvesting", ), False),
    "initial_dynamic_scale": (("initial_dynamic_scale", ), 2 ** 16),
    "load_universal_checkpoint": (("load_universal_checkpoint", ), False),
    "memory_breakdown": (("memory_breakdown", ), False),
    "mics_shard_size": (("zero_config", "mics_shard_size"), -1),
    "optimizer_legacy_fusion": (("optimizer_legacy_fusion", ), False),
    "pld_enabled": (("pld_enabled", ), False),
    "pld_params": (("pld_params", ), False),
    "sparse_gradients_enabled": (("sparse_gradients_enabled", ), False),
    "swap_tensor_config": (("swap_tensor_config", ), None),
    "use_node_local_storage": (("use_node_local_storage", ), False),
    "zero_allow_untested_optimizer": (("zero_allow_untested_optimizer", ), False),
    "zero_force_ds_cpu_optimizer": (("zero_force_ds_cpu_optimizer", ), True),
    # ---- "zero_optimization" block ---------------------------------------------------------------------------------
    "zero_allgather_bucket_size": (("zero_config", "allgather_bucket_size"), None),
    "zero_allgather_partitions": (("zero_config", "allgather_partitions"), True),
    "zero_contiguous_gradients": (("zero_config", "contiguous_gradients"), True),
    "zero_elastic_checkpoint": (("zero_config", "elastic_checkpoint"), False),
    "zero_gather_16bit_weights_on_model_save": (("zero_config", "gather_16bit_weights_on_model_save"), False),
    "zero_grad_hooks": (("zero_config", "grad_hooks"), True),
    "zero_hpz_partition_size": (("zero_config", "zero_hpz_partition_size"), 1),
    "zero_ignore_unused_parameters": (("zero_config", "ignore_unused_parameters"), True),
    "zero_legacy_stage1": (("zero_config", "legacy_stage1"), False),
    "zero_load_from_fp32_weights": (("zero_config", "load_from_fp32_weights"), True),
    "zero_log_trace_cache_warnings": (("zero_config", "log_trace_cache_warnings"), False),
    "zero_max_live_parameters": (("zero_config", "max_live_parameters"), None),
    "zero_max_reuse_distance": (("zero_config", "max_reuse_distance"), None),
    "zero_model_persistence_threshold": (("zero_config", "model_persistence_threshold"), None),
    "zero_module_granularity_threshold": (("zero_config", "module_granularity_threshold"), 0),
    "zero_multi_rank_bucket_allreduce": (("zero_config", "use_multi_rank_bucket_allreduce"), True),
    "zero_offload_optimizer": (("zero_config", "offload_optimizer"), None),
    "zero_overlap_comm": (("zero_config", "overlap_comm"), None),
    "zero_param_persistence_threshold": (("zero_config", "param_persistence_threshold"), None),
    "zero_prefetch_bucket_size": (("zero_config", "prefetch_bucket_size"), None),
    "zero_quantized_gradients": (("zero_config", "zero_quantized_gradients"), False),
    "zero_quantized_nontrainable_weights": (("zero_config", "zero_quantized_nontrainable_weights"), False),
    "zero_quantized_weights": (("zero_config", "zero_quantized_weights"), False),
    "zero_reduce_bucket_size": (("zero_config", "reduce_bucket_size"), None),
    "zero_reduce_scatter": (("zero_config", "reduce_scatter"), True),
    "zero_round_robin_gradients": (("zero_config", "round_robin_gradients"), False),
    "zero_sub_group_size": (("zero_config", "sub_group_size"), None),
    "zeropp_loco_param": (("zero_config", "zeropp_loco_param"), None),
}


def _make_getter(name: str, path: Sequence[str], default: Any) -> Callable:

    def getter(self):
        return _dig(self._config, path, default)

    getter.__name__ = name
    getter.__qualname__ = f"EngineAccessors.{name}"
    getter.__doc__ = f"``{'.'.join(path)}`` of the engine's config (default ``{default!r}``)."
    return getter


class EngineAccessors:
    """Mixed into ``DeepSpeedEngine``; expects ``_config``, ``optimizer``, ``module``, ``global_rank`` etc. of the engine."""

    # ---- getters that compute ------------------------------------------------------------------------------------------
    def autotuning_enabled(self) -> bool:
        return bool(_dig(self._config, ("autotuning_config", "enabled"), False))

    def autotuning_metric_path(self) -> str:
        path = _dig(self._config, ("autotuning_config", "metric_path"), None)
        return path or os.path.join(os.getcwd(), "autotuning_metric.json")

    def autotuning_model_info_path(self) -> str:
        path = _dig(self._config, ("autotuning_config", "model_info_path"), None)
        return path or os.path.join(os.getcwd(), "autotuning_model_info.json")

    def autotuning_profile_model_info(self) -> bool:
        return bool(self.autotuning_enabled() and _dig(self._config, ("autotuning_config", "model_info", "profile"), False))

    def flops_profiler_enabled(self) -> bool:
        return bool(_dig(self._config, ("flops_profiler_config", "enabled"), False)) or self.autotuning_enabled()

    def flops_profiler_profile_step(self) -> int:
        step = _dig(self._config, ("flops_profiler_config", "profile_step"), 1)
        return max(step, self.autotuning_start_profile_step()) if self.autotuning_enabled() else step

    def flops_profiler_detailed(self) -> bool:
        return False if self.autotuning_enabled() else bool(_dig(self._config, ("flops_profiler_config", "detailed"), True))

    def pld_theta(self):
        return _dig(self._config, ("pld_config", "theta"), _dig(self._config, ("pld_params", "theta"), 1.0))

    def pld_gamma(self):
        return _dig(self._config, ("pld_config", "gamma"), _dig(self._config, ("pld_params", "gamma"), 0.001))

    def get_pld_theta(self):
        pld = getattr(self, "progressive_layer_drop", None)
        return pld.get_theta() if pld is not None else None

    def quantize_training(self):
        """The weight-quantisation (MoQ) settings of the compression block, as the tuple the reference returns
        (``engine.py quantize_training``): in-forward flag, enabled, groups, fp16-mixed, change ratio, type, rounding,
        verbose, kernel use."""
        shared = _dig(self._config, ("compression_config", "weight_quantization", "shared_parameters"), {}) or {}
        mixed = shared.get("fp16_mixed_quantize", {}) or {}
        return (shared.get("quantize_weight_in_forward", False), shared.get("enabled", False), shared.get("quantize_groups", 1),
                mixed.get("enabled", False), mixed.get("quantize_change_ratio", 0.001), shared.get("quantization_type", "symmetric"),
                shared.get("rounding", "nearest"), shared.get("quantize_verbose", False), shared.get("quantizer_kernel", False))

    def zero_optimization_partition_gradients(self) -> bool:
        return self.zero_optimization_stage() >= 2

    def zero_optimization_partition_weights(self) -> bool:
        return self.zero_optimization_stage() >= 3

    def zero_use_cpu_optimizer(self) -> bool:
        dev = _dig(self._config, ("zero_config", "offload_optimizer", "device"), None)
        return str(getattr(dev, "value", dev)) in ("cpu", "nvme")

    def zero_partial_offload(self) -> float:
        return float(_dig(self._config, ("zero_config", "offload_optimizer", "ratio"), 1.0) or 1.0)

    def zero_nvme_offload_optimizer(self) -> bool:
        return bool(getattr(self.optimizer, "swap_optimizer", False) or getattr(self.optimizer, "optimizer_swapper", None))

    def is_first_weights_partition_group(self) -> bool:
        """True on ranks that write ZeRO-3 weight partitions: everyone without MiCS, the first shard group with it."""
        if not self.zero_optimization_partition_weights():
            return False
        shard = self.mics_shard_size()
        return shard is None or shard < 0 or self.global_rank < shard

    def is_elastic_model_parallel_supported(self) -> bool:
        if not self.elasticity_enabled():
            return False
        el = self._config.elasticity_config or {}
        mp, per_node = int(el.get("model_parallel_size", 1)), int(el.get("num_gpus_per_node", 1))
        return per_node % max(mp, 1) == 0

    def get_sequence_parallel_group(self):
        return self.seq_parallel_group

    # ---- dataset predicates / data hooks --------------------------------------------------------------------------------
    @staticmethod
    def is_map_style_dataset(obj) -> bool:
        return hasattr(obj, "__getitem__") and hasattr(obj, "__len__")

    @staticmethod
    def is_iterable_style_dataset(obj) -> bool:
        return isinstance(obj, torch.utils.data.IterableDataset)

    def set_data_post_process_func(self, post_process_func) -> None:
        if self.training_dataloader is not None:
            self.training_dataloader.post_process_func = post_process_func

    def set_custom_curriculum_learning_schedule(self, schedule_func_dict) -> None:
        loader = self.training_dataloader
        sampler = getattr(loader, "data_sampler", None) if loader is not None else None
        if sampler is not None and self.curriculum_learning_enabled():
            sampler.set_custom_curriculum_learning_schedule(schedule_func_dict)

    # ---- gradient-bucket helpers (public in the reference; Megatron-DeepSpeed calls them directly) -----------------------
    flatten = staticmethod(flatten_dense_tensors)
    unflatten = staticmethod(unflatten_dense_tensors)

    def allreduce_bucket(self, bucket, dp_group, dp_world_size=None):
        """One flat all-reduce of ``bucket`` averaged over the group (pre-/post-scaled as the config says); returns the
        flat result without touching the inputs."""
        ws = dp_world_size or dist.get_world_size(group=dp_group)
        flat = self.flatten(bucket)
        comm_dtype = self.communication_data_type() if callable(getattr(self, "communication_data_type", None)) else None
        work = flat.to(comm_dtype) if comm_dtype is not None and comm_dtype != flat.dtype else flat
        if self._config.prescale_gradients:
            pre = self.gradient_predivide_factor()
            if pre != 1.0:
                work.mul_(1.0 / pre)
            dist.all_reduce(work, group=dp_group)
            if pre != ws:
                work.mul_(pre / ws)
        else:
            dist.all_reduce(work, group=dp_group)
            work.mul_(1.0 / ws)
        if work is not flat:
            flat.copy_(work)
        return flat

    def allreduce_and_copy(self, small_bucket, dp_group, dp_world_size=None):
        reduced = self.allreduce_bucket(small_bucket, dp_group, dp_world_size)
        for buf, synced in zip(small_bucket, self.unflatten(reduced, small_bucket)):
            buf.copy_(synced)

    def allreduce_no_retain(self, bucket, dp_group, numel_per_bucket=500000000, dp_world_size=None):
        small, numel = [], 0
        for t in bucket:
            small.append(t)
            numel += t.numel()
            if numel > numel_per_bucket:
                self.allreduce_and_copy(small, dp_group, dp_world_size)
                small, numel = [], 0
        if small:
            self.allreduce_and_copy(small, dp_group, dp_world_size)

    def all_gather_scalar(self, value: torch.Tensor, dp_group):
        out = [value.new_zeros(value.size()) for _ in range(dist.get_world_size(group=dp_group))]
        dist.all_gather(out, value, group=dp_group)
        return out

    def sparse_all_gather(self, value: torch.Tensor, dp_group):
        """All-gather of tensors whose first dimension differs per rank; returns the per-rank pieces."""
        return self._all_gather_rows(value, dp_group, dist.get_world_size(group=dp_group))

    def sparse_allreduce(self, sparse, dp_group, dp_world_size=None):
        """Average a ``SparseTensor`` over the group by all-gathering (indices, values): the result holds every rank's rows."""
        ws = dp_world_size or dist.get_world_size(group=dp_group)
        orig = sparse.values.dtype
        comm_dtype = self.communication_data_type() if callable(getattr(self, "communication_data_type", None)) else None
        values = sparse.values.to(comm_dtype) if comm_dtype is not None and comm_dtype != orig else sparse.values
        if self._config.prescale_gradients:
            pre = self.gradient_predivide_factor()
            values = values * (1.0 / pre)
            post = pre / ws
        else:
            values, post = values * (1.0 / ws), 1.0
        sparse.indices = torch.cat(self.sparse_all_gather(sparse.indices, dp_group))
        gathered = torch.cat(self.sparse_all_gather(values, dp_group))
        sparse.values = (gathered * post if post != 1.0 else gathered).to(orig)
        return sparse

    def sparse_allreduce_bucket(self, bucket, dp_group, dp_world_size=None):
        return [self.sparse_allreduce(s, dp_group, dp_world_size) for s in bucket]

    def sparse_allreduce_no_retain(self, bucket, dp_group, dp_world_size=None):
        for s in self.sparse_allreduce_bucket(bucket, dp_group, dp_world_size):
            if getattr(s, "sparse_tensor", None) is not None or getattr(s, "orig_dense_tensor", None) is not None:
                dst = s.orig_dense_tensor
                dst.data = s.to_coo_tensor() if getattr(dst, "is_sparse", False) else s.to_dense().to(dst.dtype)

    # ---- ZeRO state movement -------------------------------------------------------------------------------------------
    def empty_partition_cache(self) -> None:
        """Release gathered ZeRO-3 parameters that are still resident (after an interrupted forward, or before handing
        the GPU to another model); ref: engine.py empty_partition_cache."""
        opt = self.optimizer
        if hasattr(opt, "empty_partition_cache"):
            opt.empty_partition_cache()
        elif self.zero_optimization_stage() == 3:
            from .zero.partition_parameters import ZeroParamStatus, _partition_param, is_zero_param
            for p in self.module.parameters():
                if is_zero_param(p) and p.ds_status != ZeroParamStatus.NOT_AVAILABLE and not getattr(p, "ds_persist", False):
                    _partition_param(p)
        if torch.cuda.is_available():
            torch.cuda.empty_cache()

    def offload_states(self, include=None, device="cpu", pin_memory: bool = True, non_blocking: bool = False) -> None:
        """Move the engine's optimizer state / master weights / low-precision partitions / gradients to ``device`` until
        ``reload_states``. The reference restricts this to ZeRO-3 (engine.py:3943); here it works for every stage."""
        from .zero.offload_states import offload_engine_states
        dev = str(getattr(device, "value", device))
        if dev == "none":
            return
        if dev == "nvme":
            raise ValueError("NVMe offload is not supported for offloading states.")
        if self.zero_use_cpu_optimizer():
            raise AssertionError("Moving states across devices is not supported for offloaded optimizer states.")
        offload_engine_states(self, include, torch.device(dev), pin_memory, non_blocking)

    def reload_states(self, non_blocking: bool = False) -> None:
        from .zero.offload_states import reload_engine_states
        reload_engine_states(self, non_blocking)

    # ---- loading pieces of a checkpoint -----------------------------------------------------------------------------------
    def load_module_state_dict(self, checkpoint, strict: bool = True, custom_load_fn=None, fetch_z3_params: bool = False):
        """Load ``checkpoint['module']`` into the wrapped module; with ``fetch_z3_params`` the ZeRO-3 partitions are
        gathered for the load and re-partitioned from rank 0's values after it (ref: engine.py:2790)."""
        sd = checkpoint["module"] if isinstance(checkpoint, dict) and "module" in checkpoint else checkpoint
        if getattr(self, "random_ltd_scheduler", None) is not None:
            own = self.module.state_dict().keys()
            if not any(".random_ltd_layer" in k for k in sd) and any(".random_ltd_layer" in k for k in own):
                unwrapped = {k.replace(".random_ltd_layer", ""): k for k in own}
                sd = {unwrapped.get(k, k): v for k, v in sd.items()}

        def load():
            if custom_load_fn is not None:
                custom_load_fn(src=sd, dst=self.module)
            else:
                self.module.load_state_dict(sd, strict=strict)

        if fetch_z3_params and self.zero_optimization_stage() == 3:
            from .zero.partition_parameters import GatheredParameters, is_zero_param
            params = [p for p in self.module.parameters() if is_zero_param(p)]
            with GatheredParameters(params, modifier_rank=0):
                load()
        else:
            load()
        if self.zero_optimization_stage() != 3 or fetch_z3_params:
            # the fp32 master copy follows the weights just loaded (a full checkpoint load restores it from the
            # optimizer shard afterwards anyway)
            if hasattr(self.optimizer, "refresh_fp32_params"):
                self.optimizer.refresh_fp32_params()

    @staticmethod
    def load_moe_state_dict(checkpoint_path, tag, state_dict, old_moe_load=False, model=None, mpu=None, num_experts=1,
                            checkpoint_engine=None):
        """Merge this rank's expert files ``layer_L_expert_E_mp_rank_MM_model_states.pt`` of ``checkpoint_path`` into
        ``state_dict`` under the module's own key names (ref: engine.py load_moe_state_dict, new-style layout)."""
        from ..moe.layer import MoE
        from ..utils import groups
        if old_moe_load:
            raise NotImplementedError("the pre-0.5 single-file expert layout is not written by this framework")
        mp_rank = mpu.get_model_parallel_rank() if mpu is not None else 0
        load = checkpoint_engine.load if checkpoint_engine is not None else (lambda p, map_location=None: torch.load(
            p, map_location=map_location, weights_only=False))
        layer_id = 0
        for name, m in model.named_modules():
            if not isinstance(m, MoE):
                continue
            gname = m.expert_group_name
            ep_rank = dist.get_rank(group=groups._get_expert_parallel_group(gname)) if dist.is_initialized() else 0
            n_local = m.num_local_experts
            for local_id in range(n_local):
                gid = ep_rank * n_local + local_id
                path = os.path.join(checkpoint_path, f"layer_{layer_id}_expert_{gid}_mp_rank_{mp_rank:02d}_model_states.pt")
                prefix = f"{name}.deepspeed_moe.experts.deepspeed_experts.{local_id}."
                for k, v in load(path, map_location="cpu").items():
                    state_dict[prefix + k] = v
            layer_id += 1
        return state_dict

    # ---- MoE timing, random-LTD, schedule-compiler registry -------------------------------------------------------------
    def print_forward_breakdown(self, fwd_time: float):
        """Split of the forward wall clock into gate / MoE / all-to-all time, from the MoE layers' own counters (they run
        when ``wall_clock_breakdown`` is on); ref: engine.py print_forward_breakdown."""
        from ..moe.sharded_moe import MOELayer
        from ..utils.logging import log_dist
        gate = moe = falltoall = salltoall = 0.0
        for m in self.module.modules():
            if isinstance(m, MOELayer):
                gate += m.gate_time
                moe += m.time_moe
                falltoall += m.time_falltoall
                salltoall += m.time_salltoall
        log_dist(f"time (ms) | fwd: {fwd_time:.2f} (fwd_moe: {moe:.2f}, 1st_a2a: {falltoall:.2f}, 2nd_a2a: {salltoall:.2f}, "
                 f"top_k: {gate:.2f})", ranks=[0])
        return {"fwd": fwd_time, "fwd_moe": moe, "1st_a2a": falltoall, "2nd_a2a": salltoall, "top_k": gate}

    def _configure_random_ltd_scheduler(self, configs):
        from .data_pipeline.random_ltd import RandomLTDScheduler
        return RandomLTDScheduler(configs)

    def random_ltd_initialize(self) -> None:
        """Hand the engine's random-LTD scheduler to the wrapped layers named in ``random_ltd_layer_id`` (ref:
        engine.py:691). ``convert_to_random_ltd`` calls this after wrapping."""
        from .data_pipeline.random_ltd import RandomLayerTokenDrop
        assert self.random_ltd_enabled(), "random_ltd is not enabled in data_efficiency.data_routing"
        cfg = self.random_ltd_config()
        wanted = sorted(cfg.get("random_ltd_layer_id", []))
        count = 0
        for name, layer in self.module.named_modules():
            if isinstance(layer, RandomLayerTokenDrop) and count < len(wanted) and str(wanted[count]) in name:
                layer.scheduler, layer.layer_id = self.random_ltd_scheduler, count
                count += 1
        for name, layer in self.module.named_modules():   # wrapped layers that are not in the list run every token
            if isinstance(layer, RandomLayerTokenDrop) and layer.scheduler is not self.random_ltd_scheduler:
                layer.scheduler = None
        if cfg.get("random_ltd_layer_num", count) != count:
            raise ValueError(f"random_ltd_layer_num {cfg.get('random_ltd_layer_num')} must be equivalent to the len of "
                             f"random_ltd_layer_id {count}")
        if (cfg.get("layer_token_lr_schedule") or {}).get("enabled", False):
            raise ValueError("layer_token_lr_schedule is not supported (neither is it by the reference: engine.py:707)")

    @staticmethod
    def register_compile_pass(pass_name: str, pass_fn: Callable) -> None:
        from ..compile.backend import register_compile_pass
        register_compile_pass(pass_name, pass_fn)

    @staticmethod
    def get_compile_time():
        from ..compile.backend import opt_pass_times
        return opt_pass_times


for _name, (_path, _default) in _CONFIG_GETTERS.items():
    setattr(EngineAccessors, _name, _make_getter(_name, _path, _default))
del _name, _path, _default
