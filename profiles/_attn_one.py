"""Three launches of the tensor-core attention kernel at the level-0 self-attention shape (B=64, 4096 tokens, 256 channels) and one
at the 77-key cross-attention shape -- the target of the ncu --set full capture (profiles/capture_attn.sh)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tmdiff_b200 import _lib
lib = _lib.load()
for BH, Lq, Lk, D, reps in ((64, 4096, 4096, 256, 3), (512, 4096, 77, 64, 2)):
    q = torch.randn(BH, Lq, D, device="cuda").bfloat16()
    k = torch.randn(BH, Lk, D, device="cuda").bfloat16()
    pitch = (Lk + 7) // 8 * 8
    vt = torch.randn(BH, D, pitch, device="cuda").bfloat16()
    out = torch.empty(BH, Lq, D, device="cuda")
    for _ in range(reps):
        _lib.check(lib.tmd_attention_bf16(_lib.ptr(q), _lib.ptr(k), _lib.ptr(vt), _lib.ptr(out), BH, Lq, Lk, D, pitch, D ** -0.5, _lib.stream_ptr()))
    torch.cuda.synchronize()
