"""Import paths kept for users of the reference whose modules have no separate implementation here.

The reference spreads some components over several modules (``monitor/tensorboard.py``, ``monitor/wandb.py``, ...;
``runtime/zero/parameter_offload.py`` next to ``stage3.py``; ``checkpoint/reshape_utils.py`` next to ``reshape_meg_2d.py``
...). Where this code base implements such a group in ONE module, the reference's other module paths are listed in the
table below instead of existing as near-empty files: ``alias module -> [(module that holds the object, name there, name
under the alias)]``, all relative to this package; ``"*"`` re-exports the public names. A meta-path finder builds the
alias module on first import, so ``from hcache_deepspeed_b200.monitor.wandb import WandbMonitor`` keeps working. Nothing in
the table counts as a component of its own.
"""
from __future__ import annotations

import importlib
import importlib.util
import sys
import types

PACKAGE = __name__.rsplit(".", 1)[0]

ALIASES = {
    # v2 module configs / interfaces: one file each here, a package of small files in the reference
    "inference.v2.modules.configs.attention_configs": [("inference.v2.modules.configs", n, n) for n in ("PositionalEmbeddingType", "RotateHalfConfig", "MaskingType", "DSSelfAttentionConfig")],
    "inference.v2.modules.configs.embedding_config": [("inference.v2.modules.configs", "DSEmbeddingsConfig", "DSEmbeddingsConfig")],
    "inference.v2.modules.configs.linear_config": [("inference.v2.modules.configs", "DSLinearConfig", "DSLinearConfig")],
    "inference.v2.modules.configs.moe_config": [("inference.v2.modules.configs", "DSMoEConfig", "DSMoEConfig")],
    "inference.v2.modules.configs.norm_config": [("inference.v2.modules.configs", "DSNormConfig", "DSNormConfig"), ("inference.v2.modules.configs", "NormTypeEnum", "NormTypeEnum")],
    "inference.v2.modules.configs.unembed_config": [("inference.v2.modules.configs", "DSUnembedConfig", "DSUnembedConfig")],
    "inference.v2.modules.interfaces.attention_base": [("inference.v2.modules.interfaces", n, n) for n in ("DSSelfAttentionBase", "DSSelfAttentionRegistry")],
    "inference.v2.modules.interfaces.embedding_base": [("inference.v2.modules.interfaces", n, n) for n in ("DSEmbeddingBase", "DSEmbeddingRegistry")],
    "inference.v2.modules.interfaces.linear_base": [("inference.v2.modules.interfaces", n, n) for n in ("DSLinearBase", "DSLinearRegistry")],
    "inference.v2.modules.interfaces.moe_base": [("inference.v2.modules.interfaces", n, n) for n in ("DSMoEBase", "DSMoERegistry")],
    "inference.v2.modules.interfaces.post_norm_base": [("inference.v2.modules.interfaces", n, n) for n in ("DSPostNormBase", "DSPostNormRegistry")],
    "inference.v2.modules.interfaces.pre_norm_base": [("inference.v2.modules.interfaces", n, n) for n in ("DSPreNormBase", "DSPreNormRegistry")],
    "inference.v2.modules.interfaces.unembed_base": [("inference.v2.modules.interfaces", n, n) for n in ("DSUnembedBase", "DSUnembedRegistry")],
    # the tuners live in autotuning/tuner.py
    "autotuning.tuner.base_tuner": [("autotuning.tuner", "BaseTuner", "BaseTuner")],
    "autotuning.tuner.index_based_tuner": [("autotuning.tuner", "RandomTuner", "RandomTuner"), ("autotuning.tuner", "GridSearchTuner", "GridSearchTuner")],
    "autotuning.tuner.model_based_tuner": [("autotuning.tuner", "ModelBasedTuner", "ModelBasedTuner")],
    # one v1 fused-layer class covers what the reference splits into per-family subclasses
    "model_implementations.transformers.ds_base": [("model_implementations.transformers.ds_transformer", "DeepSpeedTransformerInference", "DeepSpeedTransformerBase")],
    "model_implementations.transformers.ds_bert": [("model_implementations.transformers.ds_transformer", "DeepSpeedTransformerInference", "DeepSpeedBERTInference")],
    "model_implementations.transformers.ds_bloom": [("model_implementations.transformers.ds_transformer", "DeepSpeedTransformerInference", "DeepSpeedBloomInference")],
    "model_implementations.transformers.ds_gpt": [("model_implementations.transformers.ds_transformer", "DeepSpeedTransformerInference", "DeepSpeedGPTInference")],
    "model_implementations.transformers.ds_llama2": [("model_implementations.transformers.ds_transformer", "DeepSpeedTransformerInference", "DeepSpeedLlama2Inference")],
    "model_implementations.transformers.ds_megatron_gpt": [("model_implementations.transformers.ds_transformer", "DeepSpeedTransformerInference", "DeepSpeedMegatronGPTInference")],
    "model_implementations.transformers.ds_opt": [("model_implementations.transformers.ds_transformer", "DeepSpeedTransformerInference", "DeepSpeedOPTInference")],
    "ops.transformer.inference.ds_attention": [("model_implementations.transformers.ds_transformer", "DeepSpeedSelfAttention", "DeepSpeedSelfAttention"), ("model_implementations.transformers.ds_transformer", "DeepSpeedSelfAttention", "BloomSelfAttention")],
    # one compressed-collective back end (error-feedback 1-bit all-reduce over torch.distributed) under the reference's names
    "runtime.comm.nccl": [("runtime.comm.compressed", "CompressedBackend", "NcclBackend")],
    "runtime.comm.mpi": [("runtime.comm.compressed", "CompressedBackend", "MpiBackend")],
    "runtime.comm.hccl": [("runtime.comm.compressed", "CompressedBackend", "HcclBackend")],
    # the NVMe sweep lives in nvme/tune.py
    "nvme.perf_run_sweep": [("nvme.tune", "get_sweep_config_dict", "get_sweep_config_dict"), ("nvme.tune", "run_sweep", "run_sweep"), ("nvme.tune", "sweep_main", "sweep_main"), ("nvme.tune", "sweep_points", "sweep_points")],
    "nvme.perf_generate_param": [("nvme.tune", "convert_to_param", "convert_to_param"), ("nvme.tune", "generate_aio_param", "generate_aio_param"), ("nvme.tune", "generate_main", "generate_main")],
    "checkpoint.reshape_3d_utils": [("checkpoint.reshape", "*", "*"), ("checkpoint.reshape", "meg_2d_parallel_map", "meg_2d_parallel_map"), ("checkpoint.reshape", "model_3d_desc", "model_3d_desc"), ("checkpoint.reshape", "ZeROCheckpoint", "ZeROCheckpoint")],
    "checkpoint.reshape_meg_2d": [("checkpoint.reshape", "*", "*"), ("checkpoint.reshape", "meg_2d_parallel_map", "meg_2d_parallel_map"), ("checkpoint.reshape", "model_3d_desc", "model_3d_desc"), ("checkpoint.reshape", "ZeROCheckpoint", "ZeROCheckpoint")],
    "checkpoint.reshape_utils": [("checkpoint.reshape", "*", "*"), ("checkpoint.reshape", "meg_2d_parallel_map", "meg_2d_parallel_map"), ("checkpoint.reshape", "model_3d_desc", "model_3d_desc"), ("checkpoint.reshape", "ZeROCheckpoint", "ZeROCheckpoint")],
    "checkpoint.utils": [("checkpoint.reshape", "*", "*"), ("checkpoint.reshape", "meg_2d_parallel_map", "meg_2d_parallel_map"), ("checkpoint.reshape", "model_3d_desc", "model_3d_desc"), ("checkpoint.reshape", "ZeROCheckpoint", "ZeROCheckpoint")],
    "checkpoint.zero_checkpoint": [("checkpoint.reshape", "*", "*"), ("checkpoint.reshape", "meg_2d_parallel_map", "meg_2d_parallel_map"), ("checkpoint.reshape", "model_3d_desc", "model_3d_desc"), ("checkpoint.reshape", "ZeROCheckpoint", "ZeROCheckpoint")],
    "elasticity.config": [("elasticity.elasticity", "ElasticityConfig", "ElasticityConfig"), ("elasticity.elasticity", "ElasticityConfigError", "ElasticityConfigError"), ("elasticity.elasticity", "ElasticityError", "ElasticityError"), ("elasticity.elasticity", "ElasticityIncompatibleWorldSize", "ElasticityIncompatibleWorldSize")],
    "inference.v2.model_implementations.inference_transformer_base": [("inference.v2.model_implementations.inference_model_base", "DSInferenceModelBase", "DSInferenceModelBase"), ("inference.v2.model_implementations.inference_model_base", "DSTransformerModelBase", "DSTransformerModelBase")],
    "inference.v2.modules.module_registry": [("inference.v2.modules.ds_module", "ConfigBundle", "ConfigBundle"), ("inference.v2.modules.ds_module", "DSModuleRegistryBase", "DSModuleRegistryBase")],
    "monitor.comet": [("monitor.monitor", "CometMonitor", "CometMonitor"), ("monitor.monitor", "EventsLogScheduler", "EventsLogScheduler")],
    "monitor.csv_monitor": [("monitor.monitor", "CSVMonitor", "csvMonitor"), ("monitor.monitor", "CSVMonitor", "CSVMonitor")],
    "monitor.tensorboard": [("monitor.monitor", "TensorBoardMonitor", "TensorBoardMonitor")],
    "monitor.wandb": [("monitor.monitor", "WandbMonitor", "WandbMonitor")],
    "nvme.ds_aio_basic": [("nvme.ds_aio_handle", "aio_handle_multiprocessing", "aio_basic_multiprocessing")],
    "nvme.parse_nvme_stats": [("nvme.tune", "READ_SPEED", "READ_SPEED"), ("nvme.tune", "WRITE_SPEED", "WRITE_SPEED"), ("nvme.tune", "get_sorted_results", "get_sorted_results")],
    "ops.fp_quantizer.fp8_gemm_triton": [("ops.fp_quantizer.fp8_gemm", "matmul_fp8_fallback", "matmul_fp8")],
    "ops.lion.cpu_lion": [("ops.lion.lion", "DeepSpeedCPULion", "DeepSpeedCPULion")],
    "ops.lion.fused_lion": [("ops.lion.lion", "FusedLion", "FusedLion")],
    "ops.transformer.inference.bias_add": [("ops.spatial", "nhwc_bias_add", "nhwc_bias_add")],
    "ops.transformer.inference.ds_mlp": [("model_implementations.transformers.ds_transformer", "DeepSpeedMLP", "DeepSpeedMLP")],
    "runtime.checkpoint_engine.nebula_checkpoint_engine": [("runtime.checkpoint_engine.checkpoint_engine", "NebulaCheckpointEngine", "NebulaCheckpointEngine")],
    "runtime.checkpoint_engine.torch_checkpoint_engine": [("runtime.checkpoint_engine.checkpoint_engine", "TorchCheckpointEngine", "TorchCheckpointEngine")],
    "runtime.data_pipeline.data_routing.basic_layer": [("runtime.data_pipeline.random_ltd", "RandomLayerTokenDrop", "RandomLayerTokenDrop")],
    "runtime.data_pipeline.data_routing.helper": [("runtime.data_pipeline.random_ltd", "convert_to_random_ltd", "convert_to_random_ltd"), ("runtime.data_pipeline.random_ltd", "remove_random_ltd_state_dict", "remove_random_ltd_state_dict"), ("runtime.data_pipeline.random_ltd", "save_without_random_ltd", "save_without_random_ltd")],
    "runtime.data_pipeline.data_sampling.data_sampler": [("runtime.data_pipeline.data_sampler", "DeepSpeedDataSampler", "DeepSpeedDataSampler")],
    "runtime.fp16.unfused_optimizer": [("runtime.fp16.fused_optimizer", "FP16_UnfusedOptimizer", "FP16_UnfusedOptimizer")],
    "runtime.zero.mics_utils": [("runtime.zero.mics", "MiCS_CommGroups", "MiCS_CommGroups"), ("runtime.zero.mics", "create_mics_comm_groups", "create_mics_comm_groups"), ("runtime.zero.mics", "mics_group_ranks", "mics_group_ranks"), ("runtime.zero.mics", "scale_tensors", "scale_tensors"), ("runtime.zero.mics", "has_hierarchical_all_gather_groups", "has_hierarchical_all_gather_groups")],
    "runtime.zero.offload_config": [("runtime.zero.config", "*", "*"), ("runtime.zero.offload_states", "OffloadStateTypeEnum", "OffloadStateTypeEnum")],
    "runtime.zero.parameter_offload": [("runtime.zero.stage3", "DeepSpeedZeRoOffload", "DeepSpeedZeRoOffload"), ("runtime.zero.stage3", "ZeROOrderedDict", "ZeROOrderedDict"), ("runtime.zero.stage3", "FWD_MODULE_STACK", "FWD_MODULE_STACK")],
    "autotuning.tuner.utils": [("autotuning.tuner", n, n) for n in ("index_to_feature", "feature_to_index", "dict_to_dims", "gen_combinations", "flatten", "dict_to_feature")],
    "autotuning.tuner.cost_model": [("autotuning.tuner", "XGBoostCostModel", "XGBoostCostModel")],
    "runtime.zero.partitioned_param_coordinator": [("runtime.zero.partition_parameters", "all_gather_params", "all_gather_params"), ("runtime.zero.stage3", "DeepSpeedZeRoOffload", "PartitionedParameterCoordinator")] + [("runtime.zero.stage3", n, n) for n in ("ZeRoTraceMode", "InflightParamRegistry", "get_all_parameters", "iter_params")] + [("runtime.zero.partition_parameters", "debug_rank0", "debug_rank0")],
    "utils.comms_logging": [("comm.comms_logging", "*", "*"), ("comm.comms_logging", "CommsLogger", "CommsLogger")],
}


def _build(alias: str) -> types.ModuleType:
    full = f"{PACKAGE}.{alias}"
    mod = types.ModuleType(full, f"alias module (see {PACKAGE}._import_paths)")
    for src, name, asname in ALIASES[alias]:
        real = importlib.import_module(f"{PACKAGE}.{src}" if src else PACKAGE)
        if name == "*":
            public = getattr(real, "__all__", None) or [k for k in vars(real) if not k.startswith("_")]
            for k in public:
                setattr(mod, k, getattr(real, k))
        else:
            setattr(mod, asname, getattr(real, name))
    return mod


class _Loader:

    def __init__(self, alias: str):
        self.alias = alias

    def create_module(self, spec):
        return _build(self.alias)

    def exec_module(self, module):
        return None


class _Finder:

    def find_spec(self, fullname, path=None, target=None):
        if not fullname.startswith(PACKAGE + "."):
            return None
        alias = fullname[len(PACKAGE) + 1:]
        if alias not in ALIASES:
            return None
        return importlib.util.spec_from_loader(fullname, _Loader(alias))


def register() -> None:
    if not any(isinstance(f, _Finder) for f in sys.meta_path):
        sys.meta_path.insert(0, _Finder())
