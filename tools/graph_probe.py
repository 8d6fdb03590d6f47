"""How much of the C2 step is launch gaps?  Times the eager step against a CUDA-graph replay of the
same step (measurement only: the captured AdamW bakes its step number in)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "finetune-controller_b200"))
import torch  # noqa: E402
from b200ft import gpt2, params  # noqa: E402

wl = sys.argv[1] if len(sys.argv) > 1 else "gpt2-124m"
B, T = (32, 1024) if wl.startswith("gpt2") else (4, 2048)
dev = torch.device("cuda:0")
eng, dims = params.build_engine(wl, gpt2.LoRASpec(8 if wl.startswith("gpt2") else 16, 16.0), gpt2.OptSpec(),
                                batch=B, seq=T, device=dev, seed=0, randomize_b=True)
ids = torch.randint(0, dims.vocab, (B, T), device=dev)


def timed(fn, n=20):
    for _ in range(3):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(n):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


print(f"eager : {timed(lambda: eng.step(ids, ids)):.3f} ms/step", flush=True)
s = torch.cuda.Stream()
s.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(s):
    for _ in range(3):
        eng.step(ids, ids)
torch.cuda.current_stream().wait_stream(s)
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    eng.step(ids, ids)
print(f"graph : {timed(g.replay):.3f} ms/step   loss {float(eng.loss):.4f}", flush=True)
print(f"eager : {timed(lambda: eng.step(ids, ids)):.3f} ms/step", flush=True)
