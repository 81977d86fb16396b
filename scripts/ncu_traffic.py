"""Extract per-launch DRAM traffic and L1 wavefronts of k_network<2> from an ncu --set full report into
profiles/traffic_k_network.json (read by bench.py's roofline object).  python scripts/ncu_traffic.py rep workload n_samples"""
import csv, io, json, subprocess, sys
rep, workload, n_samples = sys.argv[1], sys.argv[2], float(sys.argv[3])
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units = rows[0], rows[1]
col = {h: i for i, h in enumerate(hdr)}
r = next(x for x in rows[2:] if 'k_network' in x[col['Kernel Name']])   # the dominant kernel's launch
def val(m, scale_unit=True):
    v = float(r[col[m]].replace(",", ""))
    u = units[col[m]]
    return v * {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1}.get(u, 1) if scale_unit else v
dram = val("dram__bytes_read.sum") + val("dram__bytes_write.sum")
cyc = val("sm__cycles_elapsed.max") if "sm__cycles_elapsed.max" in col else val("smsp__cycles_active.avg")
wf_pct = val("l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed")
d = {"kernel": r[col["Kernel Name"]].split("(")[0], "workload": workload, "samples_in_captured_launch": n_samples,
     "dram_bytes_per_launch": dram, "dram_bytes_per_sample": dram / n_samples,
     "l1_wavefronts_pct_of_peak": wf_pct, "l1_sectors_per_sample": val("l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum") / n_samples,
     "tensor_pipe_pct": val("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"),
     "duration_ms": val("gpu__time_duration.sum") / (1e6 if units[col["gpu__time_duration.sum"]] in ("ns", "nsecond") else 1e3 if units[col["gpu__time_duration.sum"]] in ("us", "usecond") else 1),
     "source": f"ncu --set full --clock-control none, one launch, {rep.split('/')[-1]}"}
json.dump(d, open("profiles/traffic_k_network.json", "w"), indent=1)
print(json.dumps(d, indent=1))
