"""Per-layer config of the v1 fused inference layer (``deepspeed/ops/transformer/inference/config.py:21-130``)."""
import json

import torch


class DeepSpeedInferenceConfig:

    def __init__(self, hidden_size=-1, intermediate_size=-1, heads=-1, num_hidden_layers=-1, layer_norm_eps=1e-12,
                 local_rank=-1, mp_size=1, dtype=torch.float16, pre_layer_norm=True, norm_type="layer_norm",
                 stochastic_mode=False, scale_attention=True, triangular_masking=True, local_attention=False, window_size=256,
                 rotary_dim=-1, rotate_half=False, rotate_every_two=True, return_tuple=True, mlp_after_attn=True,
                 mlp_act_func_type="gelu", training_mp_size=1, bigscience_bloom=False, max_out_tokens=1024,
                 min_out_tokens=1, enable_qkv_quantization=False, use_mup=False, scale_attn_by_inverse_layer_idx=False,
                 return_single_tuple=False, set_empty_params=False, transposed_mode=False, use_triton=False,
                 triton_autotune=False, num_kv=-1, rope_theta=10000, invert_mask=True, use_baddbmm=False):
        self.layer_id = -1
        self.hidden_size, self.heads, self.num_hidden_layers = hidden_size, heads, num_hidden_layers
        self.intermediate_size = intermediate_size if intermediate_size > 0 else 4 * hidden_size
        self.layer_norm_eps, self.local_rank, self.mp_size, self.dtype = layer_norm_eps, local_rank, mp_size, dtype
        self.pre_layer_norm, self.norm_type = pre_layer_norm, norm_type
        self.scale_attention, self.triangular_masking = scale_attention, triangular_masking
        self.local_attention, self.window_size = local_attention, window_size
        self.rotary_dim, self.rotate_half, self.rotate_every_two = rotary_dim, rotate_half, rotate_every_two
        self.return_tuple, self.mlp_after_attn, self.mlp_act_func_type = return_tuple, mlp_after_attn, mlp_act_func_type
        self.training_mp_size, self.bigscience_bloom = training_mp_size, bigscience_bloom
        self.max_out_tokens, self.min_out_tokens = max_out_tokens, min_out_tokens
        self.scale_attn_by_inverse_layer_idx, self.return_single_tuple = scale_attn_by_inverse_layer_idx, return_single_tuple
        self.set_empty_params, self.transposed_mode, self.use_triton = set_empty_params, transposed_mode, False
        self.num_kv = num_kv if num_kv > 0 else heads
        self.rope_theta, self.invert_mask, self.use_mup = rope_theta, invert_mask, use_mup
        self.stochastic_mode, self.enable_qkv_quantization = stochastic_mode, enable_qkv_quantization

    @classmethod
    def from_dict(cls, d):
        c = cls()
        for k, v in d.items():
            c.__dict__[k] = v
        return c

    @classmethod
    def from_json_file(cls, path):
        with open(path, "r", encoding="utf-8") as f:
            return cls.from_dict(json.loads(f.read()))
