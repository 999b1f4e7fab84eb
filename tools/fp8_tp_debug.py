"""Per-output agreement of the wf8af8 engine between tp=1 and emulated tp=2 (LoopbackWorld) on one GPU."""
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from tests.test_gpu_tp_loopback import LLAMA, _run_world  # noqa: E402

for tokens in (os.environ.get("HDS_TP_ONESHOT_TOKENS", "16"), ):
    single, multi = _run_world(LLAMA, 2, quant=True)
    names = ["prefill340", "dec1", "dec2", "dec3", "mid40", "restore_probe9"]
    for i, (w, g0, g1) in enumerate(zip(single, multi[0], multi[1])):
        cos = lambda a, b: float(torch.nn.functional.cosine_similarity(a.flatten(), b.flatten(), dim=0))  # noqa: E731
        print(f"{names[i]:16s} cos(single, r0)={cos(w, g0):.4f} cos(r0, r1)={cos(g0, g1):.4f} max|single|={float(w.abs().max()):.3f}")
