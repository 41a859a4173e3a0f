import sys; sys.path.insert(0,"tools"); sys.path.insert(0,".")
import sweep, ekat_b200
sweep.li_lm_case("f32", 1572864, 128, 128, 1, tunings=[(0,0,0),(16,0,0)])
sweep.li_lm_case("f32", 1572864, 128, 128, 10, tunings=[(0,0,0),(16,0,0)])
sweep.li_lm_case("f32", 1572864, 72, 128, 1, tunings=[(0,0,0),(16,0,0)])
for n in (3,4,5,6):
    sweep.fused_case("f64", 345600, 128, 128, n, tunings=[(0,0,3),(16,0,0)])
for n in (1,2,3):
    sweep.fused_case("f32", 345600, 128, 128, n, tunings=[(0,0,3),(32,0,0)])
sweep.fused_case("f32", 1572864, 128, 128, 10, tunings=[(0,0,3),(0,0,0)])
sweep.fused_case("f64", 1572864, 128, 128, 10, tunings=[(0,0,3),(0,0,0)])
