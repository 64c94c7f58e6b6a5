"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: one steady-state step."""
import csv, re, sys
path = sys.argv[1]; first = int(sys.argv[2]) if len(sys.argv) > 2 else 30; n = int(sys.argv[3]) if len(sys.argv) > 3 else 10
lines = [l for l in open(path) if not l.startswith("==")]
order = []
for row in csv.DictReader(lines):
    if row.get("Metric Name") != "gpu__time_duration.sum":
        continue
    v = float(row["Metric Value"].replace(",", "")); u = row["Metric Unit"]
    v = v / 1000 if u == "ns" else (v * 1000 if u == "ms" else v)
    order.append((re.sub(r"\(.*", "", row["Kernel Name"]), v, row["Grid Size"], row["Block Size"]))
step = order[first:first + n]
tot = sum(v for _, v, _, _ in step)
print("kernel,us,grid,block")
for k, v, g, b in step:
    print(f"{k},{v:.1f},{g},{b}")
print(f"TOTAL,{tot:.1f},,")
print(f"# gemm_lora_kernel share of the step: {sum(v for k, v, _, _ in step if 'gemm' in k) / tot:.3f}")
