import sys, torch
sys.path.insert(0, "/root/repo")
from pointcloudregistration_b200 import functional as F, synthetic
m = 1000000
dev = torch.device("cuda:0")
x, y = synthetic.make_pair("uniform", 1, m, m, 4)
x, y = x.to(dev), y.to(dev)
lo = (m // 3) // 8 * 8
F.nnd_forward_rows_raw(x, y, lo, lo + 125000)
torch.cuda.synchronize()
