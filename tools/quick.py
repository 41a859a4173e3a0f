import sys; sys.path.insert(0,"tools"); sys.path.insert(0,".")
import sweep, ekat_b200
for dt in ("f64","f32"):
    for n in (72, 128):
        sweep.adj_diff_case(dt, 3145728, n)
sweep.adj_diff_case("f64", 3145728, 129)
