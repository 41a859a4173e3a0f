import sys; sys.path.insert(0,"tools"); sys.path.insert(0,".")
import sweep
T=[(0,0,0),(0,0,21),(0,0,22)]
sweep.li_case("f64", 3145728, 128, 128, tunings=T)
sweep.li_case("f64", 345600, 72, 128, tunings=T)
sweep.li_case("f64", 1572864, 72, 128, tunings=T)
sweep.li_case("f64", 1572864, 128, 72, tunings=T)
sweep.li_case("f64", 1572864, 100, 100, tunings=T)
sweep.li_case("f32", 1572864, 128, 128, tunings=T)
sweep.li_case("f32", 1572864, 72, 128, tunings=T)
sweep.li_case("f32", 1572864, 128, 72, tunings=T)
