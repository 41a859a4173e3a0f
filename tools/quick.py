import sys; sys.path.insert(0,"tools"); sys.path.insert(0,".")
import sweep
for dt in ("f64","f32"):
    sweep.tridiag_case("bfb", dt, 345600, 128, tunings=[(0,0,0),(2,3,0),(4,3,0),(4,6,0)])
sweep.tridiag_case("bfb","f64",3145728,128)
sweep.tridiag_case("bfb","f64",21600,72, tunings=[(0,0,0),(2,3,0),(4,3,0)])
sweep.tridiag_case("thomas","f64",345600,128)
