"""ncu raw CSV export (`ncu -i X.ncu-rep --page raw --csv`) -> profiles/ncu_kernels.json: per kernel and launch the DRAM traffic
and the executed FP64 flops bench.py puts beside the algorithmic figures.
  usage: python tools/ncu_kernels_json.py raw.csv BATCH "source note" [out.json]
executed flops = 2 * DFMA + DADD + DMUL thread instructions (pred on) + 512 per DMMA m8n8k4 warp instruction."""
import csv, json, re, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
batch, note = int(sys.argv[2]), sys.argv[3]
out_path = sys.argv[4] if len(sys.argv) > 4 else "profiles/ncu_kernels.json"


def val(r, name):
    if name not in hdr:
        return None
    i = hdr.index(name)
    try:
        v = float(r[i].replace(",", ""))
    except ValueError:
        return None
    u = units[i].lower()
    scale = {"gbyte": 1e9, "mbyte": 1e6, "kbyte": 1e3, "byte": 1.0}.get(u, 1.0)
    return v * scale


kernels = {}
for r in rows[2:]:
    name = r[hdr.index("Kernel Name")]
    key = re.sub(r"^void\s+", "", name).split("<")[0].split("(")[0].replace("ilqrb::", "")
    rd, wr = val(r, "dram__bytes_read.sum"), val(r, "dram__bytes_write.sum")
    dfma = val(r, "smsp__sass_thread_inst_executed_op_dfma_pred_on.sum")
    dadd = val(r, "smsp__sass_thread_inst_executed_op_dadd_pred_on.sum")
    dmul = val(r, "smsp__sass_thread_inst_executed_op_dmul_pred_on.sum")
    dmma = val(r, "sm__inst_executed_pipe_tensor_subpipe_dmma.sum")
    if dmma is None:
        dmma = val(r, "smsp__inst_executed_pipe_tensor_subpipe_dmma.sum")
    e = {"kernel": name, "batch": batch, "duration_us": (val(r, "gpu__time_duration.sum") or 0) / (1e3 if units[hdr.index("gpu__time_duration.sum")] in ("nsecond", "ns") else 1.0),
         "dram_bytes": (rd or 0) + (wr or 0), "dram_bytes_read": rd, "dram_bytes_write": wr,
         "dfma_thread_inst": dfma, "dadd_thread_inst": dadd, "dmul_thread_inst": dmul, "dmma_warp_inst": dmma,
         "registers": val(r, "launch__registers_per_thread"),
         "local_ld_sectors": val(r, "l1tex__t_sectors_pipe_lsu_mem_local_op_ld.sum"), "local_st_sectors": val(r, "l1tex__t_sectors_pipe_lsu_mem_local_op_st.sum")}
    if dfma is not None:
        e["executed_flops"] = 2 * dfma + (dadd or 0) + (dmul or 0) + 512 * (dmma or 0)
    kernels.setdefault(key, e)  # first captured launch of each kernel
try:
    old = json.load(open(out_path))
except Exception:
    old = {"kernels": {}}
old["kernels"].update(kernels)
old["source"] = note
json.dump(old, open(out_path, "w"), indent=1)
print(json.dumps(kernels, indent=1))
