"""Dev tool: a small pass through every kernel of libnnd_b200.so with every tensor in its own cudaMalloc
(PYTORCH_NO_CUDA_MEMORY_CACHING=1, so an out-of-bounds write lands outside torch's pooled blocks and faults instead of
silently hitting a neighbour).  Written as a compute-sanitizer target (memcheck / racecheck / initcheck / synccheck);
compute-sanitizer is closed on this GPU pool, so it runs bare and cross-checks the three forward paths against
each other (they must agree bit for bit).  Shapes are small on purpose."""
import os
import sys
from pathlib import Path

os.environ.setdefault("PYTORCH_NO_CUDA_MEMORY_CACHING", "1")  # every tensor its own cudaMalloc: memcheck sees real bounds

import torch  # noqa: E402

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pointcloudregistration_b200 import _cabi, synthetic  # noqa: E402
from pointcloudregistration_b200 import functional as F  # noqa: E402
from pointcloudregistration_b200 import pytorch3d_compat as P3D  # noqa: E402

dev = torch.device("cuda:0")
lib = _cabi.load()


def same(a, b, what):
    for x, y in zip(a, b):
        assert torch.equal(x, y), what


def say(msg):
    print("[selfcheck]", msg, flush=True)


# small clouds: the single-launch kernels (default) and, with that path switched off, the two-pass kernel in all three
# work shapes + the one-direction entry point + both backward paths
for small_max in (-1, 0):
    keep = lib.nnd_set_small_max_points(small_max)
    for b, n, m in ((2, 300, 700), (1, 2048, 1024), (40, 64, 70), (1, 5000, 4100)):
        x1, x2 = (t.to(dev) for t in synthetic.make_pair("crosscopy", b, n, m, 7))
        d1, d2, i1, i2 = F.nnd_forward_raw(x1, x2)
        dq, iq = F.nnd_query_raw(x1, x2)
        same((d1, i1), (dq, iq), "query != forward")
        gd1, gd2 = (t.to(dev) for t in synthetic.grad_pair(b, n, m, 7))
        F.nnd_backward_raw(x1, x2, gd1, gd2, i1, i2)
        torch.cuda.synchronize()
        say(f"{'single-launch' if lib.nnd_forward_kind(b, n, m) == 0 else 'two-pass'} b={b} n={n} m={m} ok")
    lib.nnd_set_small_max_points(keep)
# every configuration of the single-launch forward, and a ragged batch (lengths read by the kernels)
x1, x2 = (t.to(dev) for t in synthetic.make_pair("lattice", 3, 700, 900, 9))
ref_small = F.nnd_forward_raw(x1, x2)
for w in (2, 4, 8, 16):
    for q in (1, 2, 4, 8):
        for t in (1, 2, 4, 8, 16):
            if w % t == 0:
                lib.nnd_set_small_config(w, q, t)
                same(F.nnd_forward_raw(x1, x2), ref_small, f"small config {w},{q},{t}")
lib.nnd_set_small_config(0, 0, 0)
l1, l2 = torch.tensor([700, 13, 0], device=dev), torch.tensor([900, 1, 5], device=dev)
rd = F.nnd_forward_ragged_raw(x1, x2, l1, l2)
gd1, gd2 = (t.to(dev) for t in synthetic.grad_pair(3, 700, 900, 9))
F.nnd_backward_ragged_raw(x1, x2, gd1, gd2, rd[2], rd[3], l1, l2)
F.nnd_knn_backward_ragged_raw(x1, x2, gd1, rd[2], l1, l2)
torch.cuda.synchronize()
say("single-launch configurations + ragged batch ok")
keep_small = lib.nnd_set_small_max_points(0)  # the sections below are about the general paths

# persistent symmetric kernel, then the same problem through the panel kernel (1, 2, 4 tiles per panel) and through
# row ranges (normal and role-swapped): all bit-identical
b, n, m = 1, 8200, 9001
x1, x2 = (t.to(dev) for t in synthetic.make_pair("clustered", b, n, m, 8))
before = lib.nnd_set_panel_min_rows(1 << 40)
ref = F.nnd_forward_raw(x1, x2)
torch.cuda.synchronize()
say("symmetric persistent ok")
lib.nnd_set_panel_min_rows(1)
for tiles in (4, 2, 1):
    lib.nnd_set_panel_tiles(tiles)
    same(F.nnd_forward_raw(x1, x2), ref, f"panel scan ({tiles} tiles) != persistent scan")
    torch.cuda.synchronize()
    say(f"panel scan, {tiles} tile(s) per panel ok")
lib.nnd_set_panel_tiles(0)
g = F.row_group()
cuts = [0, 2048 // g * g, 6000 // g * g, n]
keys = None
rows_d, rows_i = [], []
for lo, hi in zip(cuts[:-1], cuts[1:]):
    d, i, k = F.nnd_forward_rows_raw(x1, x2, lo, hi)
    say(f"rows [{lo},{hi}) scan kind {lib.nnd_rows_scan_kind(1, n, m, lo, hi)} ok")
    rows_d.append(d)
    rows_i.append(i)
    keys = k if keys is None else torch.minimum(keys, k)
d2, i2 = F.nnd_finalize_cols_raw(x1, x2, keys)
same((torch.cat(rows_d, 1), d2, torch.cat(rows_i, 1), i2), ref, "merged row ranges != single launch")
lib.nnd_set_panel_min_rows(before)

# small clouds through the row-range entry point (mid / small shapes of the symmetric kernel)
for b, n, m in ((3, 64, 500), (2, 700, 300)):
    x1, x2 = (t.to(dev) for t in synthetic.make_pair("lattice", b, n, m, 9))
    ref = F.nnd_forward_raw(x1, x2)
    d, i, k = F.nnd_forward_rows_raw(x1, x2, 0, n)
    d2, i2 = F.nnd_finalize_cols_raw(x1, x2, k)
    same((d, d2, i, i2), ref, "symmetric small shape != two-pass")
    torch.cuda.synchronize()
    say(f"symmetric small shape b={b} n={n} m={m} ok")

# fused chamfer epilogue + backward through autograd (the NDP call)
a = synthetic.make_pair("uniform", 1, 3000, 2900, 10)[0].to(dev).requires_grad_(True)
c = synthetic.make_pair("uniform", 1, 3000, 2900, 10)[1].to(dev)
loss = P3D.compute_truncated_chamfer_distance(a, c, trunc=0.05)
loss.backward()
torch.cuda.synchronize()
say(f"fused chamfer epilogue ok, loss {loss.item():.6e}")
say("ALL OK")
lib.nnd_set_small_max_points(keep_small)
