import sys; sys.path.insert(0,"tools"); sys.path.insert(0,".")
import sweep
sweep.li_case("f64", 21600, 72, 128, 1, tunings=[(0, 0, 0), (8, 0, 0), (16,0,0)])
sweep.li_case("f64", 21600, 72, 128, 16)
sweep.li_case("f64", 345600, 72, 128, 1, tunings=[(0, 0, 0), (8, 0, 0), (16, 0, 0), (24, 0, 0)])
sweep.li_case("f64", 345600, 128, 128, 1, tunings=[(0, 0, 0), (8,0,0), (16, 0, 0)])
sweep.li_case("f32", 345600, 72, 128, 1, tunings=[(0, 0, 0), (8,0,0), (16, 0, 0), (32,0,0)])
sweep.li_case("f64", 345600, 72, 128, 1, nf=4)
sweep.li_case("f64", 3145728, 128, 128, 1)
