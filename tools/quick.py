import sys; sys.path.insert(0,"tools"); sys.path.insert(0,".")
import sweep, ekat_b200
V=[(0,0,0),(0,0,10),(0,0,11)]
for nlev in (32,64,72,96,128,160,256):
    sweep.tridiag_case("bfb","f64",345600*128//nlev,nlev, tunings=V)
for nlev in (72,128,256):
    sweep.tridiag_case("bfb","f32",345600*128//nlev,nlev, tunings=V)
    sweep.tridiag_case("bfb","f64",345600*128//nlev,nlev, layout=ekat_b200.COL_MAJOR, tunings=V)
    sweep.tridiag_case("bfb","f32",345600*128//nlev,nlev, layout=ekat_b200.COL_MAJOR, tunings=V)
