"""Dev tool (GPU box): the reuse path of the pruned search against the from-scratch path, step by step on a moving cloud."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pointcloudregistration_b200 import ndp_loop, synthetic  # noqa: E402

dev = torch.device("cuda:0")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
x0, y = synthetic.make_pair("uniform", 1, n, n, 3)
x0, y = x0.to(dev), y.to(dev)
a = ndp_loop.ChamferLoop(1, n, n, dev, search="pruned", resort_every=0, use_graph=False)
b = ndp_loop.ChamferLoop(1, n, n, dev, search="pruned", resort_every=8, use_graph=False)
a.set_target(y)
b.set_target(y)
x = x0.clone()
for it in range(60):
    la, ga = a.step(x)
    lb, gb = b.step(x)
    same = [bool(torch.equal(p, q)) for p, q in ((a.dist1, b.dist1), (a.dist2, b.dist2), (a.idx1, b.idx1), (a.idx2, b.idx2))]
    gsame = bool(torch.equal(ga, gb))
    if not all(same) or not gsame or it % 10 == 0:
        print(it, "forward equal:", same, "grad equal:", gsame, "loss", float(la), float(lb), flush=True)
    if not all(same):
        bad = (a.idx1 != b.idx1).nonzero()
        print("first idx1 mismatches", bad[:5].tolist(), a.idx1[0, bad[:5, 1]].tolist(), b.idx1[0, bad[:5, 1]].tolist(),
              a.dist1[0, bad[:5, 1]].tolist(), b.dist1[0, bad[:5, 1]].tolist())
        break
    x = x - 0.01 * n * ga
