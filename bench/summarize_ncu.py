"""Turn an `.ncu-rep` (ncu --set full) into the compact table committed under profiles/.
usage: python bench/summarize_ncu.py gpurun_out/prof_hot.ncu-rep > profiles/<name>.md"""
import csv, io, json, os, subprocess, sys

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
peaks = {}
pp = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
if os.path.isfile(pp):
    peaks = json.load(open(pp))
M = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
     "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
     "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
     "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
     "launch__block_size", "launch__shared_mem_per_block_dynamic"]
ix = {m: hdr.index(m) for m in M if m in hdr}
kn = hdr.index("Kernel Name")
print(f"# ncu --set full summary of `{os.path.basename(rep)}`\n")
print("Measured peaks used for the fractions (MEASURED_PEAKS.json): "
      f"HBM copy {peaks.get('hbm_gbs', 'n/a')} GB/s, cuBLAS bf16 {peaks.get('bf16_tflops', 'n/a')} TFLOP/s burst.\n")
print("| kernel | grid x block | regs | dyn smem | time us | DRAM rd+wr MB | achieved GB/s (of measured) | DRAM % (ncu) | tensor pipe % | warps active % |")
print("|---|---|---|---|---|---|---|---|---|---|")
seen = set()
for r in data:
    name = r[kn].split("(")[0].replace("void ", "").replace("<unnamed>::", "")
    key = (name, r[ix["launch__grid_size"]])
    if key in seen:
        continue
    seen.add(key)
    def f(m, d=1.0):
        try:
            return float(r[ix[m]].replace(",", "")) * d
        except Exception:
            return float("nan")
    t = f("gpu__time_duration.sum")
    tu = units[ix["gpu__time_duration.sum"]]
    t_us = t / 1000 if tu == "ns" else (t * 1000 if tu == "ms" else t)
    def mb(m):
        v, u = f(m), units[ix[m]].lower()
        return v * {"byte": 1e-6, "kbyte": 1e-3, "mbyte": 1, "gbyte": 1e3}.get(u, 1)
    traffic = mb("dram__bytes_read.sum") + mb("dram__bytes_write.sum")
    gbs = traffic / 1e3 / (t_us / 1e6) if t_us > 0 else 0
    frac = f" ({gbs / peaks['hbm_gbs']:.2f})" if peaks.get("hbm_gbs") else ""
    print(f"| `{name}` | {r[ix['launch__grid_size']]} x {r[ix['launch__block_size']]} | {r[ix['launch__registers_per_thread']]} | "
          f"{r[ix['launch__shared_mem_per_block_dynamic']]} | {t_us:.1f} | {traffic:.1f} | {gbs:.0f}{frac} | "
          f"{f('gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed'):.1f} | "
          f"{f('sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active'):.1f} | "
          f"{f('sm__warps_active.avg.pct_of_peak_sustained_active'):.1f} |")
