"""Raw tcgen05.mma rate: cycles per (128 x 256 x 16 per SM) UMMA, single CTA vs CTA pair, vs commit frequency."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from diffsinger_b200 import _native as nat  # noqa: E402

L = nat.lib()
for pair in (0, 1):
    for grid in (148,):
        for commit_every in (0, 1, 2, 3, 7):
            out = torch.zeros(2 * grid, dtype=torch.int64, device="cuda")
            n = 4096
            for _ in range(2):
                nat.check(L.ds_debug_mma_rate(pair, grid, n, commit_every, out.data_ptr(), 0))
            torch.cuda.synchronize()
            o = out.view(grid, 2)[:: (2 if pair else 1)].float()
            print(f"pair={pair} grid={grid:3d} mode={commit_every} (1=commit/4mma 2=wait 4=desc): issue {o[:, 0].mean() / n:7.1f} cyc/mma, "
                  f"total {o[:, 1].mean() / n:7.1f} cyc/mma (max {o[:, 1].max() / n:7.1f})")
