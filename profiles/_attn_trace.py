"""TMD_ATTN_TRACE=1: per-key-block cycle split of the softmax warp (attn_umma_kernel, long-sequence variant)."""
import os, sys, torch
os.environ["TMD_ATTN_TRACE"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tmdiff_b200 import _lib
lib = _lib.load()
for BH, Lq, Lk, D in ((64, 4096, 4096, 256), (64, 4096, 4096, 128), (64, 1024, 1024, 512)):
    q = torch.randn(BH, Lq, D, device="cuda").bfloat16()
    k = torch.randn(BH, Lk, D, device="cuda").bfloat16()
    vt = torch.randn(BH, D, Lk, device="cuda").bfloat16()
    out = torch.empty(BH, Lq, D, device="cuda")
    for _ in range(2):
        _lib.check(lib.tmd_attention_bf16(_lib.ptr(q), _lib.ptr(k), _lib.ptr(vt), _lib.ptr(out), BH, Lq, Lk, D, Lk, D ** -0.5, _lib.stream_ptr()))
    torch.cuda.synchronize()
