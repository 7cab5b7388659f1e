import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from kair_b200 import ops
dev = "cuda"; B = 8; T = B * 128 * 128
qkv = (torch.randn(T, 576, device=dev) * 0.5).bfloat16(); out = torch.empty(T, 192, device=dev, dtype=torch.bfloat16)
table = torch.randn(225, 6, device=dev) * 0.1
l = ops.build_window_attn_fwd(qkv, table, out, batch=B, H=128, W=128, num_heads=6, window=8, shift=4)
os.environ.pop("KB_AF_TS", None)
for _ in range(3): l.run(ops.stream_handle())
torch.cuda.synchronize()
os.environ["KB_AF_TS"] = "1"
l.run(ops.stream_handle()); torch.cuda.synchronize()
