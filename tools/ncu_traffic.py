"""Summarise an `ncu --set full` report: per kernel (name up to the template arguments) launches, time,
DRAM bytes, L2->SM bytes, tensor-pipe activity.  Usage: python tools/ncu_traffic.py report.ncu-rep|raw_page.csv [key]
With a key (e.g. c3_b1024_tf32) the per-kernel DRAM bytes per launch are merged into
profiles/r2_traffic.json, which bench.py reads for `roofline.traffic`."""
import collections, csv, io, json, os, re, subprocess, sys

rep = sys.argv[1]
raw = open(rep).read() if rep.endswith(".csv") else \
    subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[0]
col = {h: i for i, h in enumerate(hdr)}
def num(r, name):
    v = r[col[name]].replace(",", "") if name in col and r[col[name]] else "0"
    try:
        return float(v)
    except ValueError:
        return 0.0
units = dict(zip(hdr, rows[1]))
def to_bytes(r, name):
    u = units.get(name, "")
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
    return num(r, name) * scale
def to_us(r, name):
    u = units.get(name, "")
    scale = {"ns": 1e-3, "us": 1, "ms": 1e3, "s": 1e6}.get(u, 1)
    return num(r, name) * scale
agg = collections.OrderedDict()
for r in rows[2:]:
    if len(r) < len(hdr):
        continue
    name = re.sub(r"\(.*", "", r[col["Kernel Name"]]).replace("void ", "").replace("gsc::", "")
    a = agg.setdefault(name, dict(n=0, us=0.0, rd=0.0, wr=0.0, l2sm=0.0, tensor=0.0, ipc=0.0))
    a["n"] += 1
    a["us"] += to_us(r, "gpu__time_duration.sum")
    a["rd"] += to_bytes(r, "dram__bytes_read.sum")
    a["wr"] += to_bytes(r, "dram__bytes_write.sum")
    a["l2sm"] += to_bytes(r, "l1tex__m_xbar2l1tex_read_bytes.sum")
    a["tensor"] += num(r, "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active")
    a["ipc"] += num(r, "smsp__issue_active.avg.pct_of_peak_sustained_active")
print("| kernel | launches | us total | DRAM read MB | DRAM written MB | DRAM GB/s | L2->SM MB | tensor pipe % | issue active % |")
print("|---|---|---|---|---|---|---|---|---|")
for k, a in agg.items():
    gbs = (a["rd"] + a["wr"]) / (a["us"] * 1e-6) / 1e9 if a["us"] else 0
    print(f"| `{k}` | {a['n']} | {a['us']:.1f} | {a['rd']/1e6:.1f} | {a['wr']/1e6:.1f} | {gbs:.0f} | {a['l2sm']/1e6:.1f} | "
          f"{a['tensor']/a['n']:.1f} | {a['ipc']/a['n']:.1f} |")
if len(sys.argv) > 2:
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "r2_traffic.json")
    try:
        table = json.load(open(path))
    except (OSError, ValueError):
        table = {}
    entry = table.setdefault(sys.argv[2], {})
    for k, a in agg.items():
        base = k.split("<")[0]
        e = entry.setdefault(base, dict(launches=0, dram_bytes=0.0, us=0.0))
        e["launches"] += a["n"]; e["dram_bytes"] += a["rd"] + a["wr"]; e["us"] += a["us"]
    for e in entry.values():
        e["dram_bytes_per_launch"] = e["dram_bytes"] / max(e["launches"], 1)
    json.dump(table, open(path, "w"), indent=1, sort_keys=True)
    print("merged into", path)
