"""Text summary of an ncu --set full report in the format of profiles/ncu_forest_r01.txt: the headline counters per kernel
(raw page) and the stall reasons + hottest SASS lines (source page).
    python tools/ncu_summary.py gpurun_out/x.ncu-rep "command line it was captured with" > profiles/ncu_x_r02.txt"""
import csv, io, subprocess, sys

rep, cmd = sys.argv[1], (sys.argv[2] if len(sys.argv) > 2 else "")
WANT = ["gpu__time_duration.sum", "sm__cycles_elapsed.avg", "sm__cycles_elapsed.sum", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_subpipe_imma_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_tmem.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "launch__grid_size", "launch__block_size", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "launch__waves_per_multiprocessor"]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw))); hdr, units = rows[0], rows[1]; idx = {h: i for i, h in enumerate(hdr)}
print(f"# ncu --set full --clock-control none --import-source on   ({cmd})")
print("# (captured after the same command exited 0 without ncu; per-launch times under ncu are cold-cache and serialised)")
for r in rows[2:]:
    print(f"\n## {r[idx['Kernel Name']]}")
    for w in WANT:
        if w in idx and r[idx[w]] not in ("", "n/a"): print(f"{w},{r[idx[w]]},{units[idx[w]]}")
    try:
        wf = float(r[idx["l1tex__data_pipe_lsu_wavefronts_mem_shared.sum"]].replace(",", "")); cyc = float(r[idx["sm__cycles_elapsed.sum"]].replace(",", ""))
        if wf: print(f"derived: shared-memory pipe busy = wavefronts / SM-cycles = {wf / cyc:.3f}")
    except Exception: pass
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src))); blocks, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name": cur = {"name": r[1], "hdr": None, "data": []}; blocks.append(cur); continue
    if cur is None: continue
    if r and r[0] == "Address": cur["hdr"] = r; continue
    if cur["hdr"] and len(r) >= len(cur["hdr"]) - 2: cur["data"].append(r)
for b in blocks[:1]:
    H = {h: i for i, h in enumerate(b["hdr"])}; data = b["data"]
    g = lambda r, k: int(float(r[H[k]] or 0)) if k in H and H[k] < len(r) and r[H[k]] not in ("", "-") else 0
    stalls = [h for h in b["hdr"] if h.startswith("stall_") and "Not Issued" not in h]
    ns = sum(g(r, "# Samples") for r in data) or 1
    print(f"\n# stall reasons + hottest SASS lines ({b['name'][:90]})\nsamples {ns} warp-instructions {sum(g(r, 'Instructions Executed') for r in data)}")
    for s, v in sorted(((s, sum(g(r, s) for r in data)) for s in stalls), key=lambda kv: -kv[1])[:8]: print(f"  {s:28s} {v:8d} {v / ns:.3f}")
    for r in sorted(data, key=lambda r: -g(r, "# Samples"))[:14]:
        print(f'{g(r, "# Samples"):6d} {g(r, "Instructions Executed"):9d} {r[H["Source"]][:66]:66s} lsb {g(r, "stall_long_sb"):5d} ssb {g(r, "stall_short_sb"):5d} math {g(r, "stall_math"):5d} wait {g(r, "stall_wait"):5d} bar {g(r, "stall_barrier"):5d}')
