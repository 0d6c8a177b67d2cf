"""Time the attention backward (3B tp1 shape) once; used to compare experimental builds (NXD_B200_LIB=...)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from neuronx_distributed_llama3_2_b200.ops import attention as A
B, Hq, Hkv, S, D = 1, 24, 8, 8192, int(os.environ.get("D", 128))
q = torch.randn(B, Hq, S, D, device="cuda").bfloat16().requires_grad_()
k = torch.randn(B, Hkv, S, D, device="cuda").bfloat16().requires_grad_()
v = torch.randn(B, Hkv, S, D, device="cuda").bfloat16().requires_grad_()
do = torch.randn(B, Hq, S, D, device="cuda").bfloat16()
def t(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
tf = t(lambda: A.flash_attn_func(q, k, v, causal=True))
def fb():
    o = A.flash_attn_func(q, k, v, causal=True); o.backward(do); q.grad = k.grad = v.grad = None
tfb = t(fb)
fl = 4.0 * B * Hq * S * S * D * 0.5
print(f"{os.environ.get('NXD_B200_LIB', 'default'):45s} fwd {tf:.3f} ms {fl/tf/1e9:6.0f} TF   bwd {tfb-tf:.3f} ms {2.5*fl/(tfb-tf)/1e9:6.0f} TF")
