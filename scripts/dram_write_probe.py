"""Probe: how does ncu's dram__bytes_write count plain coalesced stores on this GPU?  (288 MB fill / copy)"""
import torch
n = 36_000_000
a = torch.empty(n, dtype=torch.float64, device="cuda")
b = torch.ones(n, dtype=torch.float64, device="cuda")
torch.cuda.synchronize()
for _ in range(3):
    a.fill_(2.0)
    a.copy_(b)
    c = a * 2.0
torch.cuda.synchronize()
print("done")
