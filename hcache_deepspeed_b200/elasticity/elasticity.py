"""Elastic batch-size arithmetic (ref: deepspeed/elasticity/elasticity.py:83-300, config.py, constants.py).

Given the acceptable micro-batch sizes and a maximum global batch size, find the global batch size that is
divisible by the largest number of GPU counts in ``[min_gpus, max_gpus]`` — so a job can lose or gain GPUs and
resume with the same effective batch. v0.1 works per GPU, v0.2 per node with a model-parallel degree.
"""
from __future__ import annotations

import json
import math
import os
from typing import List, Optional, Tuple

ELASTICITY = "elasticity"
LATEST_ELASTICITY_VERSION = 0.2
MINIMUM_DEEPSPEED_VERSION = "0.3.8"
HCN_LIST = [1, 2, 4, 6, 12, 24, 36, 48, 60, 120, 180, 240, 360, 720, 840, 1260, 1680, 2520, 5040, 7560, 10080, 15120,
            20160, 25200, 27720, 45360, 50400, 55440, 83160, 110880, 166320, 221760, 277200, 332640, 498960, 554400,
            665280, 720720]


class ElasticityError(Exception):
    pass


class ElasticityConfigError(ElasticityError):
    pass


class ElasticityIncompatibleWorldSize(ElasticityError):
    pass


class ElasticityConfig:

    def __init__(self, param_dict: dict):
        self.enabled = param_dict.get("enabled", False)
        if self.enabled:
            if "max_train_batch_size" not in param_dict:
                raise ElasticityConfigError("Elasticity config missing max_train_batch_size")
            if "micro_batch_sizes" not in param_dict:
                raise ElasticityConfigError("Elasticity config missing micro_batch_sizes")
        self.max_acceptable_batch_size = param_dict.get("max_train_batch_size", 2000)
        self.micro_batches = param_dict.get("micro_batch_sizes", [2, 4, 6])
        if not isinstance(self.micro_batches, list) or not all(isinstance(m, int) and m > 0 for m in self.micro_batches):
            raise ElasticityConfigError(f"micro_batch_sizes must be a list of positive ints, got {self.micro_batches}")
        self.min_gpus = param_dict.get("min_gpus", 1)
        self.max_gpus = param_dict.get("max_gpus", 10000)
        if self.min_gpus < 1 or self.max_gpus < 1 or self.max_gpus < self.min_gpus:
            raise ElasticityConfigError("invalid min_gpus / max_gpus")
        self.model_parallel_size = param_dict.get("model_parallel_size", 1)
        self.num_gpus_per_node = param_dict.get("num_gpus_per_node", 1)
        self.min_time = param_dict.get("min_time", 0)
        self.version = param_dict.get("version", 0.2)
        self.prefer_larger_batch_size = param_dict.get("prefer_larger_batch", True)
        self.ignore_non_elastic_batch_info = param_dict.get("ignore_non_elastic_batch_info", False)


def elasticity_enabled(ds_config: dict) -> bool:
    return ds_config.get(ELASTICITY, {}).get("enabled", False)


def get_candidate_batch_sizes(base_list: List[int], max_acceptable_batch_size: int) -> List[int]:
    candidates = set()
    for base in base_list:
        if base >= max_acceptable_batch_size:
            candidates.add(base)
            continue
        for hcn in HCN_LIST:
            if base * hcn > max_acceptable_batch_size:
                break
            candidates.add(base * hcn)
    return sorted(candidates)


def get_valid_gpus(batch_size: int, micro_batches: List[int], min_valid_gpus: int, max_valid_gpus: int) -> List[int]:
    valid = set()
    for mb in micro_batches:
        if batch_size % mb:
            continue
        max_gpus = batch_size // mb
        for i in range(1, int(math.isqrt(max_gpus)) + 1):
            if max_gpus % i == 0:
                for g in (i, max_gpus // i):
                    if min_valid_gpus <= g <= max_valid_gpus:
                        valid.add(g)
    return sorted(valid)


def get_best_candidates(candidates: List[int], micro_batches: List[int], min_gpus: int, max_gpus: int,
                        prefer_larger: bool) -> Tuple[int, List[int]]:
    best_count, best_bs, best_gpus = 0, micro_batches[0], []
    for bs in candidates:
        gpus = get_valid_gpus(bs, micro_batches, min_gpus, max_gpus)
        if len(gpus) > best_count or (len(gpus) == best_count and ((prefer_larger and bs > best_bs) or
                                                                     (not prefer_larger and bs < best_bs))):
            best_count, best_bs, best_gpus = len(gpus), bs, gpus
    return best_bs, best_gpus


def get_compatible_gpus(micro_batches: List[int], max_acceptable_batch_size: int, min_gpus: Optional[int] = None,
                        max_gpus: Optional[int] = None, prefer_larger: bool = True) -> Tuple[int, List[int]]:
    """v0.1: best global batch size and the GPU counts it supports."""
    min_gpus = min_gpus or 1
    max_gpus = max_gpus or max_acceptable_batch_size // min(micro_batches)
    if any(mb > max_acceptable_batch_size for mb in micro_batches):
        raise ValueError(f"All micro batches must be <= max_acceptable_batch_size {max_acceptable_batch_size}")
    lcm = 1
    for mb in micro_batches:
        lcm = lcm * mb // math.gcd(lcm, mb)
    base_list = sorted(set(micro_batches + [lcm]))
    candidates = get_candidate_batch_sizes(base_list, max_acceptable_batch_size)
    return get_best_candidates(candidates, micro_batches, min_gpus, max_gpus, prefer_larger)


def _get_compatible_gpus_v02(micro_batches, max_acceptable_batch_size, current_num_gpus, min_gpus=None, max_gpus=None,
                             prefer_larger=True, num_gpus_per_node=1, model_parallel_size=1):
    if num_gpus_per_node % model_parallel_size != 0:
        raise ElasticityError(f"num_gpus_per_node {num_gpus_per_node} must be divisible by model_parallel_size "
                              f"{model_parallel_size}")
    dp_per_node = num_gpus_per_node // model_parallel_size

    def pick_microbatch(final_bs):
        cand = [mb for mb in micro_batches if (final_bs // (current_num_gpus // model_parallel_size)) % mb == 0]
        return (max(cand) if prefer_larger else min(cand)) if cand else None

    final_bs, valid_nodes = get_compatible_gpus(micro_batches, int(max_acceptable_batch_size / dp_per_node),
                                                int((min_gpus or 1) / num_gpus_per_node) or 1,
                                                int((max_gpus or 10000) / num_gpus_per_node) or 1, prefer_larger)
    final_bs = int(final_bs) * dp_per_node
    valid_dp = [n * dp_per_node for n in valid_nodes]
    if current_num_gpus // model_parallel_size in valid_dp:
        return final_bs, valid_dp, pick_microbatch(final_bs)
    current_dp = (current_num_gpus / num_gpus_per_node) * dp_per_node
    cands = []
    for mb in micro_batches:
        bs = int(max_acceptable_batch_size // (mb * current_dp)) * mb * current_dp
        if bs > 0:
            cands.append(int(bs))
    if not cands:
        raise ElasticityIncompatibleWorldSize(f"no valid batch size for {current_num_gpus} GPUs")
    final_bs = max(cands) if prefer_larger else min(cands)
    return final_bs, [int(current_dp)], pick_microbatch(final_bs)


def compute_elastic_config(ds_config: dict, target_deepspeed_version: str = "0.16.8", world_size: int = 0,
                           return_microbatch: bool = False):
    """Returns ``(final_batch_size, valid_gpus[, micro_batch_size])`` (ref: elasticity.py:233-300)."""
    if not isinstance(ds_config, dict):
        raise ValueError(f"Expected ds_config to be a dictionary but received a {type(ds_config)}")
    if ELASTICITY not in ds_config:
        raise ElasticityError(f"'{ELASTICITY}' is missing from config json, please add it if running an elastic job.")
    if not ds_config[ELASTICITY].get("enabled", False):
        raise ElasticityError("Elasticity is disabled, please enable it ('enabled':true) if running an elastic job.")
    cfg = ElasticityConfig(ds_config[ELASTICITY])
    if float(cfg.version) > LATEST_ELASTICITY_VERSION:
        raise ElasticityConfigError(f"Attempting to run elasticity version {cfg.version} but runtime only supports up "
                                    f"to {LATEST_ELASTICITY_VERSION}")
    if cfg.model_parallel_size > 1 and float(cfg.version) != 0.2:
        raise ElasticityConfigError("Elasticity v0.1 does not support model-parallel training")
    micro_batch = None
    if float(cfg.version) == 0.1:
        final_bs, valid_gpus = get_compatible_gpus(cfg.micro_batches, cfg.max_acceptable_batch_size, cfg.min_gpus,
                                                   cfg.max_gpus, cfg.prefer_larger_batch_size)
        final_bs = int(final_bs)
    else:
        current = world_size if world_size > 0 else int(os.environ.get("WORLD_SIZE", 0))
        if current <= 0:
            raise ElasticityConfigError("Elasticity v0.2 needs WORLD_SIZE (or world_size=) to compute a valid batch size")
        final_bs, valid_gpus, micro_batch = _get_compatible_gpus_v02(
            cfg.micro_batches, cfg.max_acceptable_batch_size, current, cfg.min_gpus, cfg.max_gpus,
            cfg.prefer_larger_batch_size, cfg.num_gpus_per_node, cfg.model_parallel_size)
        final_bs = int(final_bs)
    if world_size > 0 and float(cfg.version) == 0.1:
        if world_size not in valid_gpus:
            raise ElasticityIncompatibleWorldSize(f"World size ({world_size}) is not valid with the current list of valid "
                                                  f"GPU counts: {valid_gpus}")
        for mb in sorted(set(cfg.micro_batches), reverse=True):
            if final_bs // world_size % mb == 0:
                micro_batch = mb
                break
    if return_microbatch:
        return final_bs, valid_gpus, micro_batch
    return final_bs, valid_gpus


def ensure_immutable_elastic_config(runtime_elastic_config_dict: dict) -> None:
    """The resource scheduler publishes the elastic config it planned with in ``DEEPSPEED_ELASTICITY_CONFIG``; a runtime
    config that disagrees on batch size, micro batches or version is an error (ref: elasticity.py:208-231)."""
    import json
    from ..utils.logging import logger
    from .constants import DEEPSPEED_ELASTICITY_CONFIG
    raw = os.environ.get(DEEPSPEED_ELASTICITY_CONFIG)
    if raw is None:
        logger.warning("DEEPSPEED_ELASTICITY_CONFIG is not set: cannot check that the resource scheduler scales this job over "
                       "compatible GPU counts")
        return
    sched, run = ElasticityConfig(json.loads(raw)), ElasticityConfig(runtime_elastic_config_dict)
    for field in ("max_acceptable_batch_size", "micro_batches", "version"):
        if getattr(sched, field) != getattr(run, field):
            raise ElasticityConfigError(f"elastic config seen by the resource scheduler ({field}={getattr(sched, field)}) does not "
                                        f"match the runtime config ({field}={getattr(run, field)})")


def is_torch_elastic_compatible() -> bool:
    """torch.distributed.elastic (the agent ``elasticity/elastic_agent.py`` builds on) exists since torch 1.11."""
    import torch
    try:
        major, minor = (int(x) for x in torch.__version__.split(".")[:2])
    except ValueError:
        return True
    return (major, minor) >= (1, 11)
