"""Dev tool: ndp_loop.ChamferLoop per-iteration time with and without the warm start, graph and eager."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pointcloudregistration_b200 import ndp_loop, synthetic  # noqa: E402

dev = torch.device("cuda:0")
for n in (100_000, 262_144):
    x0, y = (t.to(dev) for t in synthetic.make_pair("uniform", 1, n, n, 3))
    for use_graph in (True, False):
        for warm in (False, True):
            loop = ndp_loop.ChamferLoop(1, n, n, dev, use_graph=use_graph, warm=warm)
            loop.set_target(y)
            x = x0.clone()
            loop.step(x)
            torch.cuda.synchronize()
            marks = []
            e0 = torch.cuda.Event(enable_timing=True)
            e0.record()
            for it in range(100):
                _, g = loop.step(x)
                x = x - 0.01 * n * g
                if it % 25 == 24:
                    e = torch.cuda.Event(enable_timing=True)
                    e.record()
                    marks.append(e)
            torch.cuda.synchronize()
            prev, parts = e0, []
            for e in marks:
                parts.append(prev.elapsed_time(e) / 25)
                prev = e
            print(f"N=M={n} graph={use_graph} warm={warm}: ms/iteration over iterations 0-24, 25-49, 50-74, 75-99: "
                  + ", ".join(f"{p:.3f}" for p in parts) + f"  loss {loop.loss_value():.3e}", flush=True)
