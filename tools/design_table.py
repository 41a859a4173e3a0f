"""Markdown rows of DESIGN.md section 4's table from a bench line:  python tools/design_table.py profiles/rXX_bench.json"""
import json
import sys

j = json.loads(open(sys.argv[1]).read().splitlines()[-1])
r = j["roofline"]
print("| Kernel | Case | ms | GB/s | of measured peak |")
print("|---|---|---|---|---|")
print(f"| `{r['kernel']}` | bench step: thomas f64, 1 rhs, 3,145,728 x 128 | {r['avg_launch_ms']:.3f} | {r['achieved']:.0f} | **{r['frac']:.3f}** |")
s = r["second_kernel"]
print(f"| `{s['kernel']}` | bench step: LinInterp f64 128->128, 3,145,728 columns | {s['avg_launch_ms']:.3f} | {s['achieved']:.0f} | **{s['frac']:.3f}** |")
for k in j["other_kernels"]:
    print(f"| `{k['kernel']}` | {k['case']} | {k['ms']:.4g} | {k['gbs']:.0f} | {k['frac']:.3f} |")
