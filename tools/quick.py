import sys; sys.path.insert(0,"tools"); sys.path.insert(0,".")
import sweep, ekat_b200
for dt in ("f64","f32"):
    for lay in (ekat_b200.COL_MAJOR, ekat_b200.LEV_MAJOR):
        for ser in (True, False):
            sweep.reduction_case(dt, 3145728, 128, layout=lay, serialize=ser)
sweep.reduction_case("f64", 345600, 128)
sweep.reduction_case("f64", 345600, 72)
