import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ekat_b200._lib import lib
L = lib(); dev = torch.device("cuda:0"); out = torch.zeros(1, dtype=torch.int64, device=dev)
for dt in ("f64", "f32"):
    fn = getattr(L, f"ekat_b200_selftest_div_{dt}")
    for mode in (2, 3, 4, 5, 6, 7):
        out.zero_(); fn(1 << 26, 999 + mode, mode, out.data_ptr(), None); torch.cuda.synchronize()
        print(dt, "mode", mode, int(out.item()), "of", 1 << 27, flush=True)
