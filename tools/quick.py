import sys; sys.path.insert(0,"tools"); sys.path.insert(0,".")
import sweep, ekat_b200
T=[(16,0,0),(0,0,3)]
sweep.fused_case("f64", 1572864, 128, 128, 10, tunings=T)
sweep.fused_case("f32", 1572864, 128, 128, 10, tunings=T)
sweep.fused_case("f64", 345600, 72, 72, 10, tunings=T)
sweep.fused_case("f64", 345600, 128, 128, 4, tunings=T)
sweep.fused_case("f64", 345600, 128, 128, 1, tunings=T)
