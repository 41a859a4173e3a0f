import sys; sys.path.insert(0,"tools"); sys.path.insert(0,".")
import sweep, ekat_b200
sweep.tridiag_case("bfb","f64",1572864,128, nrhs=10, tunings=[(0,0,0),(0,0,13)])
sweep.tridiag_case("bfb","f64",345600,128, nrhs=10, tunings=[(0,0,0),(0,0,13)])
sweep.tridiag_case("bfb","f32",345600,128, nrhs=10, tunings=[(0,0,0),(0,0,13)])
sweep.tridiag_case("bfb","f64",345600,128, nrhs=3, tunings=[(0,0,0),(0,0,13)])
