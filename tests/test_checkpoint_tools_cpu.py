"""zero_to_fp32 consolidation, universal checkpoint conversion + resume under another dp degree, checkpoint inspection
(reference: tests/unit/checkpoint/{test_zero_optimizer,test_universal_checkpoint}.py)."""
import os

import pytest
import torch

from tests.dist_utils import run_distributed
from tests.runtime_helper import ckpt_tools_worker

pytestmark = pytest.mark.timeout(600)


@pytest.mark.parametrize("stage", [1, 2, 3])
def test_zero_to_fp32_consolidates_shards(stage, tmp_path):
    from hcache_deepspeed_b200.checkpoint import (DeepSpeedCheckpoint, convert_zero_checkpoint_to_fp32_state_dict,
                                                  get_fp32_state_dict_from_zero_checkpoint)
    d = str(tmp_path / "ck")
    states = run_distributed(ckpt_tools_worker, 2, stage, d, "save")
    sd = get_fp32_state_dict_from_zero_checkpoint(d)          # uses the 'latest' file
    for k, v in states[0].items():
        torch.testing.assert_close(sd[k], v, rtol=1e-6, atol=1e-7)
    out = convert_zero_checkpoint_to_fp32_state_dict(d, str(tmp_path / "fp32"), tag="t3")
    assert os.path.exists(out) and set(torch.load(out, weights_only=False)) >= set(states[0])
    ck = DeepSpeedCheckpoint(os.path.join(d, "t3"))
    assert ck.original_dp_degree == 2 and ck.original_tp_degree == 1 and len(ck.zero_files) == 2


def test_universal_checkpoint_resumes_with_other_dp_degree(tmp_path):
    from hcache_deepspeed_b200.checkpoint import convert_to_universal
    d = str(tmp_path / "ck")
    states = run_distributed(ckpt_tools_worker, 2, 2, d, "save")
    uni = convert_to_universal(d, str(tmp_path / "ck" / "t3_universal"), tag="t3")
    assert os.path.isdir(os.path.join(uni, "zero", "head.weight")) and os.path.exists(os.path.join(uni, "zero", "head.weight", "exp_avg.pt"))
    res = run_distributed(ckpt_tools_worker, 1, 1, uni, "universal")   # dp 2 / stage 2  ->  dp 1 / stage 1
    assert res[0]["restored"] == len(states[0]) and res[0]["exp_avg_norm"] > 0
    for k, v in states[0].items():
        torch.testing.assert_close(res[0]["state"][k], v, rtol=1e-6, atol=1e-7)
