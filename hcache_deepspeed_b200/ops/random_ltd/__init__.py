from .dropping_utils import GatherTokens, ScatterTokens, bert_sample_tokens, gpt_sample_tokens  # noqa: F401
