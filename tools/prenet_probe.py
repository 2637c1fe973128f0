"""Run the masking + prenet fc1 kernel a few times on the bench shape ([1024, 800, 80] -> 256) -- the target of
`ncu -k regex:prenet_fc1` captures; prints the CUDA-event time of the fc1 kernel alone (zero fill: no mean pass)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from unast_b200.engine import get_engine

eng = get_engine()
B, T, N = 1024, 800, 256
x = torch.rand((B, T, 80), device="cuda")
spans = torch.tensor([[10, 30, 400, 50]] * B, dtype=torch.int32, device="cuda")
keep = (torch.rand((B, T), device="cuda") > 0.1).float()
packed, n_out = eng.prenet_pack(torch.randn(N, 80, device="cuda") / 9, torch.randn(N, device="cuda"))
y = torch.empty((B, T, N), device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for zero in (True, False):
    ts = []
    for i in range(6):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        eng.mel_prenet_fc1(x, packed, n_out, spans=spans, keep=keep, replace_with_zero=zero, out=y)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    print("replace_with_zero" if zero else "mean fill (two kernels)", "ms", ["%.4f" % t for t in ts[2:]])
eng.check()
