"""dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernels, from `ncu --set full` raw CSV pages,
into profiles/ncu_traffic.json (what bench.py reports as roofline.traffic, so the number always comes from a committed
capture and names it).

    ncu -i gpurun_out/r2_full_c2.ncu-rep --page raw --csv > /tmp/raw_c2.csv
    python tools/ncu_traffic.py c2/words=/tmp/raw_c2.csv c5/words=/tmp/raw_c5.csv > profiles/ncu_traffic.json
"""
import csv
import json
import sys

OPS = (("crc32c_chunks", "crc32c"), ("snappy_compress", "compress"), ("snappy_decompress", "decompress"),
       ("snappy_seg_", "decompress"), ("snappy_gather_kernel", "pack"))  # an op of several kernels: their sum


def to_bytes(value, unit):
    v = float(value.replace(",", ""))
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}[unit]


out = {"_source": "ncu --set full --clock-control none captures of tools/devbench.py, one launch of every kernel of the op (profiles/r2_ncu_full_key_metrics_*.txt)"}
for spec in sys.argv[1:]:
    key, path = spec.split("=")
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    ix = {h: i for i, h in enumerate(hdr)}
    for r in rows[2:]:
        if len(r) < len(hdr):
            continue
        name = r[ix["Kernel Name"]]
        for pat, op in OPS:
            if pat in name:
                rd = to_bytes(r[ix["dram__bytes_read.sum"]], units[ix["dram__bytes_read.sum"]])
                wr = to_bytes(r[ix["dram__bytes_write.sum"]], units[ix["dram__bytes_write.sum"]])
                out[f"{key}/{op}"] = out.get(f"{key}/{op}", 0) + int(rd + wr)
                break
print(json.dumps(out, indent=1))
