"""Which fp32 arithmetic does torch use on the GPU for `tensor / python_float`?  (bin index of train/losses/cdf_loss.py:138)"""
import numpy as np
import torch
bw = (15.0 - 0.0) / 250
x = torch.rand(20_000_000, device="cuda") * 16.0
edges = torch.arange(0, 260, device="cuda", dtype=torch.float32) * np.float32(bw)
x = torch.cat([x, edges, torch.nextafter(edges, edges + 1), torch.nextafter(edges, edges - 1)])
ref = ((x - 0.0) / bw)
cands = {"x * f32(1/f32(bw))": x * torch.tensor(np.float32(1.0) / np.float32(bw), device="cuda"),
         "x * f32(1/bw)": x * torch.tensor(np.float32(1.0 / bw), device="cuda"),
         "x / f32(bw)": x / torch.tensor(np.float32(bw), device="cuda"),
         "f32(f64(x) / bw)": (x.double() / bw).float()}
for k, v in cands.items():
    print(k, "value mismatches", int((v != ref).sum()), "bin mismatches", int((v.long() != ref.long()).sum()))
ref2 = ((x - 0.0) / bw + 0.5)
print("x=0.06 ->", float(((torch.tensor([0.06], device="cuda") - 0.0) / bw)[0]))
