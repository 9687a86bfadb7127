"""Three forwards of the ResNet-50 assessor at P particles (ncu target: 54 launches per forward)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "gan-discovery-pso_b200"))
import torch
import bench
from psob import engine

P = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
net = engine.AssessorNet(bench.make_assessor().state_dict(), P)
img = torch.rand(P, 784, device="cuda")
for _ in range(3):
    out = net.forward(img, rescale=True)
torch.cuda.synchronize()
print("logits", tuple(out.shape), float(out.abs().mean()))
