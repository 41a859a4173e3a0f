"""profiles/traffic.json from an `ncu --set full` capture of bench.py's two timed kernels:

    ncu --set full --clock-control none --import-source on -k 'regex:k_tridiag_tma_shared|k_lin_interp' -s 6 -c 2 \
        -f -o gpurun_out/rNN_full_bench python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu --no-extras --no-strong
    python tools/make_traffic.py gpurun_out/rNN_full_bench.ncu-rep "<command above>" [out.json]

bench.py attach_traffic() uses an entry only for the same kernel name, the same column count and an unchanged source file
(sha256 of the .cu), so a stale capture can never decorate a newer kernel."""
import csv
import hashlib
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402  (csrc_hash / file_hash: the very functions attach_traffic compares with)

# demangled-name fragment -> (key in traffic.json, ekat_b200_last_kernel() name, source file)
KERNELS = [("k_tridiag_tma_shared", "tridiag_thomas_f64", "tridiag_tma_shared<f64,thomas>", "ekat_b200/csrc/tridiag_tma_shared.cu"),
           ("k_lin_interp", "lin_interp_f64", "lin_interp<f64,fused>", "ekat_b200/csrc/lin_interp.cu")]


def main():
    rep, source = sys.argv[1], sys.argv[2]
    out_path = sys.argv[3] if len(sys.argv) > 3 else os.path.join(ROOT, "profiles", "traffic.json")
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(txt.splitlines()))
    hdr, data = rows[0], rows[2:]
    col = {n: hdr.index(n) for n in ("Kernel Name", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__time_duration.sum",
                                     "launch__grid_size")}
    units = rows[1]

    def to_bytes(v, unit):
        scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}[unit]
        return float(v.replace(",", "")) * scale

    kernels = {}
    for d in data:
        name = d[col["Kernel Name"]]
        for frag, key, api_name, rel in KERNELS:
            if frag in name and key not in kernels:
                rd = to_bytes(d[col["dram__bytes_read.sum"]], units[col["dram__bytes_read.sum"]])
                wr = to_bytes(d[col["dram__bytes_write.sum"]], units[col["dram__bytes_write.sum"]])
                kernels[key] = {"kernel": api_name, "ncu_kernel_name": name[:160], "file": rel, "file_sha16": bench.file_hash(rel),
                                "dram_bytes": int(rd + wr), "dram_read_bytes": int(rd), "dram_write_bytes": int(wr),
                                "ncu_duration": d[col["gpu__time_duration.sum"]] + " " + units[col["gpu__time_duration.sum"]],
                                "grid": d[col["launch__grid_size"]]}
    tj = {"ncol": 3145728, "nlev": 128, "source": source, "csrc_sha16": bench.csrc_hash(), "kernels": kernels,
          "report_sha16": hashlib.sha256(open(rep, "rb").read()).hexdigest()[:16]}
    with open(out_path, "w") as f:
        json.dump(tj, f, indent=1)
    print(json.dumps(tj, indent=1))


if __name__ == "__main__":
    main()
