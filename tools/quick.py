import sys, os; sys.path.insert(0,"tools"); sys.path.insert(0,".")
import sweep, ekat_b200
for mb in (0, 24, 48, 64, 96, 128, 192, 256, 512):
    os.environ["EKAT_B200_SLAB_MB"] = str(mb)
    print("slab MB", mb, flush=True)
    sweep.fused_case("f64", 1572864, 128, 128, 10)
