import sys; sys.path.insert(0,"tools"); sys.path.insert(0,".")
import sweep, ekat_b200
T=[(0,0,0),(16,0,0),(0,0,3)]
sweep.fused_case("f64", 1572864, 128, 128, 10, tunings=T[:1])
sweep.fused_case("f32", 1572864, 128, 128, 10, tunings=T)
sweep.fused_case("f32", 345600, 72, 72, 10, tunings=T)
sweep.fused_case("f32", 345600, 128, 128, 4, tunings=T)
sweep.fused_case("f64", 345600, 128, 128, 2, tunings=T[1:])
sweep.fused_case("f64", 345600, 128, 128, 3, tunings=T[1:])
sweep.fused_case("f64", 345600, 100, 60, 5, tunings=T[1:])
