"""Per-launch role timings (TMD_UMMA_TRACE=1) of one denoiser evaluation at the bench workload."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import wavbest_ref
from oracle.inputs import make_inputs
from tmdiff_b200 import WavBEST
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
m = WavBEST(channels=[32, 64, 128, 256], dtype="bf16", text_embeddings=wavbest_ref.seeded_text_embeddings(), max_batch=B)
m.load_state_dict(wavbest_ref.seeded_weights((32, 64, 128, 256)))
m = m.cuda()
pan, ms, x = [t.cuda() for t in make_inputs(B, 4, 64, 64)]
t = torch.full((B,), 500.0).cuda()
m(x, t, PAN=pan, MS=ms, prompt="QB")
torch.cuda.synchronize()
os.environ["TMD_UMMA_TRACE"] = os.environ.get("TRACE_LEVEL", "1")
m(x, t, PAN=pan, MS=ms, prompt="QB")
torch.cuda.synchronize()
