"""Can the UNMODIFIED reference serve Llama through its v1 kernel-injection path on this box?
Step 1: JIT-build its `transformer_inference` op for sm_100 (InferenceBuilder().load(), the reference's own build path);
step 2: deepspeed.init_inference(replace_with_kernel_inject=True) over a small HF Llama and one forward.
Everything (build log included) goes to stdout; artefacts are copied to gpurun_out/ref_ext/ so a successful build can be
kept with the tree. Evidence for BASELINE.md section 3 / 4."""
import glob
import os
import shutil
import sys
import time
import traceback

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "baseline", "_ref"))
ext_dir = os.path.join(ROOT, "baseline", "_ref_ext")
os.makedirs(ext_dir, exist_ok=True)
os.environ["TORCH_EXTENSIONS_DIR"] = ext_dir
os.environ.setdefault("TORCH_CUDA_ARCH_LIST", "10.0")
os.environ["DS_SKIP_CUDA_CHECK"] = "1"
os.environ["DS_ACCELERATOR"] = "cuda"
os.environ.setdefault("MAX_JOBS", "16")
t0 = time.time()
ok_build = False
try:
    import torch
    import deepspeed
    from deepspeed.ops.op_builder import InferenceBuilder
    b = InferenceBuilder()
    print("InferenceBuilder.is_compatible:", b.is_compatible(verbose=True), flush=True)
    mod = b.load(verbose=True)
    print(f"BUILD OK in {time.time() - t0:.0f} s: {mod}", flush=True)
    ok_build = True
except Exception:
    traceback.print_exc()
    print(f"BUILD FAILED after {time.time() - t0:.0f} s", flush=True)
out = os.path.join(ROOT, "gpurun_out", "ref_ext")
os.makedirs(out, exist_ok=True)
for f in glob.glob(os.path.join(ext_dir, "**", "*.so"), recursive=True) + glob.glob(os.path.join(ext_dir, "**", "build.ninja"), recursive=True):
    shutil.copy(f, out)
    print("kept", f, os.path.getsize(f), "bytes")
if ok_build:
    try:
        from transformers import LlamaConfig, LlamaForCausalLM
        cfg = LlamaConfig(vocab_size=32000, hidden_size=4096, intermediate_size=11008, num_hidden_layers=2, num_attention_heads=32,
                          num_key_value_heads=32, max_position_embeddings=4096)
        torch.set_default_dtype(torch.bfloat16)
        with torch.device("cuda"):
            model = LlamaForCausalLM(cfg).eval()
        torch.set_default_dtype(torch.float32)
        ref_ids = torch.randint(0, 32000, (1, 64), device="cuda")
        with torch.no_grad():
            want = model(input_ids=ref_ids).logits.float()
        eng = deepspeed.init_inference(model, dtype=torch.bfloat16, replace_with_kernel_inject=True)
        with torch.no_grad():
            got = eng(input_ids=ref_ids).logits.float()
        print("KERNEL-INJECT FORWARD OK; max abs diff vs HF eager:", (got - want).abs().max().item(), flush=True)
    except Exception:
        traceback.print_exc()
        print("KERNEL-INJECT FAILED", flush=True)
