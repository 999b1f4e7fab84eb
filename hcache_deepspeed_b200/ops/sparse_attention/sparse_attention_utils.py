"""Helpers to retrofit block-sparse attention into a pretrained encoder
(``ops/sparse_attention/sparse_attention_utils.py:13-208``): longer position tables, padding to the block size."""
from __future__ import annotations

import torch
from torch.nn import functional as F

from .bert_sparse_self_attention import BertSparseSelfAttention
from .sparsity_config import SparsityConfig


class SparseAttentionUtils:

    @staticmethod
    def extend_position_embedding(model, max_position: int):
        """Tile the learned position table up to ``max_position`` (BERT: whole table; RoBERTa: rows after the 2 reserved)."""
        if hasattr(model, "bert"):
            emb = model.bert.embeddings.position_embeddings
            orig = emb.weight.shape[0]
            assert max_position > orig
            reps = (max_position + orig - 1) // orig
            emb.weight.data = emb.weight.data.repeat(reps, 1)[:max_position].clone()
            emb.num_embeddings = max_position
            model.bert.embeddings.position_ids = torch.arange(max_position).unsqueeze(0)
        elif hasattr(model, "roberta"):
            emb = model.roberta.embeddings.position_embeddings
            orig = emb.weight.shape[0] - 2
            max_position += 2
            reps = (max_position - 2 + orig - 1) // orig
            body = emb.weight.data[2:].repeat(reps, 1)[:max_position - 2]
            emb.weight.data = torch.cat([emb.weight.data[:2], body]).clone()
            emb.num_embeddings = max_position
            model.roberta.embeddings.position_ids = torch.arange(max_position).unsqueeze(0)
        else:
            raise ValueError("Please extend 'extend_position_embedding' function to support your model type. It "
                             "currently only supports 'bert' & 'roberta'!")
        model.config.max_position_embeddings = max_position
        return model

    @staticmethod
    def update_tokenizer_model_max_length(tokenizer, max_position: int):
        tokenizer.model_max_length = max_position
        if hasattr(tokenizer, "init_kwargs"):
            tokenizer.init_kwargs["model_max_length"] = max_position
        return tokenizer

    @staticmethod
    def replace_model_self_attention_with_sparse_self_attention(model, max_position: int,
                                                                sparsity_config=SparsityConfig(num_heads=4)):
        if hasattr(model, "bert"):
            model.config.max_position_embeddings = max_position
            SparseAttentionUtils.replace_self_attention_layer_with_sparse_self_attention_layer(
                model.config, model.bert.encoder.layer, sparsity_config)
        elif hasattr(model, "roberta"):
            model.config.max_position_embeddings = max_position + 2
            SparseAttentionUtils.replace_self_attention_layer_with_sparse_self_attention_layer(
                model.config, model.roberta.encoder.layer, sparsity_config)
        else:
            raise ValueError("Please extend 'update_model_self_attention_to_sparse_self_attention' function to support "
                             "your model type. It currently only supports 'bert' & 'roberta'!")
        return model

    @staticmethod
    def replace_self_attention_layer_with_sparse_self_attention_layer(config, layers, sparsity_config=SparsityConfig(num_heads=4)):
        for layer in layers:
            new = BertSparseSelfAttention(config, sparsity_config)
            old = layer.attention.self
            for name in ("query", "key", "value"):
                getattr(new, name).load_state_dict(getattr(old, name).state_dict())
            layer.attention.self = new
        return layers

    @staticmethod
    def pad_to_block_size(block_size: int, input_ids, attention_mask, token_type_ids, position_ids, inputs_embeds,
                          pad_token_id: int, model_embeddings):
        """Right-pad every given input to a multiple of ``block_size``; returns ``(pad_len, *padded inputs)``."""
        seq_len = (input_ids if input_ids is not None else inputs_embeds).shape[1]
        pad_len = (block_size - seq_len % block_size) % block_size
        if pad_len > 0:
            if inputs_embeds is not None:
                pad_ids = inputs_embeds.new_full((inputs_embeds.shape[0], pad_len), pad_token_id, dtype=torch.long)
                inputs_embeds = torch.cat([inputs_embeds, model_embeddings(pad_ids)], dim=-2)
            if input_ids is not None:
                input_ids = F.pad(input_ids, (0, pad_len), value=pad_token_id)
            if position_ids is not None:
                position_ids = F.pad(position_ids, (0, pad_len), value=pad_token_id)
            if attention_mask is not None:
                attention_mask = F.pad(attention_mask, (0, pad_len), value=False)
            if token_type_ids is not None:
                token_type_ids = F.pad(token_type_ids, (0, pad_len), value=0)
        return pad_len, input_ids, attention_mask, token_type_ids, position_ids, inputs_embeds

    @staticmethod
    def unpad_sequence_output(pad_len: int, sequence_output):
        return sequence_output[:, :-pad_len] if pad_len > 0 else sequence_output
