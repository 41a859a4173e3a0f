"""Scratch driver for one-off A/B runs on the GPU box: edit, then `gpurun -- python tools/quick.py`."""
import sys; sys.path.insert(0, "tools"); sys.path.insert(0, ".")
import sweep
sweep.fused_case("f64", 1572864, 128, 128, 10, tunings=[(0, 0, 0), (0, 0, 3)])
