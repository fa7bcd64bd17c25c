import sys, torch
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pointcloudregistration_b200 import functional as F, synthetic
kind = sys.argv[1] if len(sys.argv) > 1 else "uniform"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 1000000
b = int(sys.argv[3]) if len(sys.argv) > 3 else 1
dev = torch.device("cuda:0")
x, y = synthetic.make_pair(kind, b, n, n, 4)
x, y = x.to(dev), y.to(dev)
F.nnd_forward_pruned_raw(x, y)
torch.cuda.synchronize()
