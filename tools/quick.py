import sys; sys.path.insert(0,"tools"); sys.path.insert(0,".")
import sweep, ekat_b200
sweep.li_lm_case("f64", 1572864, 128, 128, 10, tunings=[(0,0,0),(0,0,1)])
sweep.li_lm_case("f64", 1572864, 128, 128, 1, tunings=[(0,0,0),(0,0,2)])
sweep.li_lm_case("f64", 1572864, 128, 128, 2, tunings=[(0,0,0),(0,0,2)])
sweep.li_lm_case("f64", 1572864, 128, 128, 3, tunings=[(0,0,0),(0,0,1)])
sweep.li_lm_case("f32", 1572864, 128, 128, 10, tunings=[(0,0,0),(0,0,1)])
sweep.fused_case("f64", 1572864, 128, 128, 10)
