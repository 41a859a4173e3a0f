"""Scratch driver for one-off A/B runs on the GPU box: edit, then `gpurun -- python tools/quick.py`."""
import sys; sys.path.insert(0, "tools"); sys.path.insert(0, ".")
import sweep
T = [(0, 0, 0), (0, 1, 0)]
sweep.fused_case("f64", 1572864, 128, 128, 10, tunings=T)
sweep.fused_case("f32", 1572864, 128, 128, 10, tunings=T)
sweep.fused_case("f64", 345600, 72, 72, 10, tunings=T)
sweep.fused_case("f64", 345600, 128, 128, 5, tunings=T)
