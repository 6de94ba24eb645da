"""DRAM traffic per launch of each kernel class from `ncu --set full` reports of tools/prof_case.py.

    python tools/ncu_traffic.py <clips in the capture>[:<clips of the captured log-mel launch>] gpurun_out/x_gemm.ncu-rep ... > profiles/traffic.json

Writes {class: {"bytes_per_launch_per_clip": ..., "launches": n, "clips": c, "source": [...]}}; bench.py scales it
by the clips of its own launches to fill roofline.traffic (dram__bytes_read.sum + dram__bytes_write.sum)."""
import csv, io, json, subprocess, sys

CLASS = {"gemm_bf16_kernel": "gemm_tcgen05", "attention_bf16_kernel": "attention_tcgen05", "attention3_bf16_kernel": "attention_tcgen05", "logmel_kernel": "logmel",
         "layernorm_kernel": "layernorm", "normalize_apply_kernel": "normalize"}
# the log-mel kernel is persistent (1-D grid), so the clips of the captured launch are given: prof_case.py runs it
# over max(clips, 128) clips in launches of <= 52 (one L2-resident chunk), and the capture takes a full one
clips, _, logmel_clips = sys.argv[1].partition(":")
clips, logmel_clips = int(clips), int(logmel_clips or 52)
out = {}
for path in sys.argv[2:]:
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(hdr)}
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    for r in data:
        name = next((c for k, c in CLASS.items() if k in r[idx["Kernel Name"]]), None)
        if name is None:
            continue
        total = sum(float(r[idx[m]]) * scale[units[idx[m]]] for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
        n_clips = logmel_clips if name == "logmel" else clips
        e = out.setdefault(name, {"bytes": 0.0, "launches": 0, "clips": n_clips, "source": []})
        e["bytes"] += total
        e["launches"] += 1
        if path not in e["source"]:
            e["source"].append(path)
for e in out.values():
    e["bytes_per_launch_per_clip"] = round(e.pop("bytes") / e["launches"] / e["clips"], 1)
print(json.dumps(out, indent=1))
