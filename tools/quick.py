import sys; sys.path.insert(0,"tools"); sys.path.insert(0,".")
import sweep, ekat_b200
V=[(0,0,0),(0,0,12),(2,0,0),(4,0,0)]
sweep.tridiag_case("thomas","f32",345600,128, tunings=V)
sweep.tridiag_case("thomas","f32",3145728,128, tunings=V)
sweep.tridiag_case("thomas","f32",345600,72, tunings=V)
sweep.tridiag_case("thomas","f64",345600,128, nA=4, nrhs=4, tunings=V)
sweep.tridiag_case("thomas","f32",345600,128, layout=ekat_b200.COL_MAJOR, tunings=[(0,0,0),(0,0,12)])
