"""The results tables of DESIGN.md §6 from bench.py lines: python scripts/results_table.py profiles/r02_bench_n*_head.json"""
import json, sys

rows = []
for f in sys.argv[1:]:
    for line in open(f):
        if line.startswith("{"):
            rows.append(json.loads(line))
rows.sort(key=lambda d: d["n_gpus"])
print("| N | `value` GB/s (÷ D2H sum) | e2e take GB/s (÷ ceiling) | restore GB/s (÷ ceiling) | `async_take` blocked ms (median) | ceilings of that box: D2H per GPU; tmpfs write / read GB/s; cores |")
print("|---|---|---|---|---|---|")
for d in rows:
    r, c = d["roofline"], d["ceilings"]
    print(f"| {d['n_gpus']} | {d['value']:.1f} ({r['host_link']['frac']:.2f}) | {d['e2e']['value']:.1f} ({r['e2e_take']['frac']:.2f}) | "
          f"{d['e2e']['restore_value']:.1f} ({r['restore']['frac']:.2f}) | {d['async_take_blocked_note']['median_ms']:.1f} | "
          f"{c['d2h_per_gpu_gbps']}; {c['tmpfs_write_gbps']} / {c['tmpfs_read_gbps']}; {c['host_cores']} |")
print()
print("| N | C3 tensors / GB per rank: take, restore, blocked ms | C4 shape, take at N, reshard-restore at N/2 | C5 shape: take N, restore N/2, take N/2, restore N | row → column at N (NVLink in, per GPU) | `read_object` 1 M rows |")
print("|---|---|---|---|---|---|")
for d in rows:
    s = d["suite"]; c3, c4, c5 = s["c3"], s["c4_reshard"], s["c5"]
    rc = c5.get("row_to_col")
    rc_txt = "-" if not rc else (f"{rc['gbps']} GB/s ({rc['frac']}); {rc['nvlink_gb']} GB over NVLink at {rc['nvlink_gbps_per_gpu_in']:.0f} GB/s = "
                                 f"{rc['nvlink_frac_of_900']} of 900")
    print(f"| {d['n_gpus']} | {c3['tensors_per_rank']} / {c3['gb_per_rank']}: {c3['take_gbps']} ({c3['take_frac']}), {c3['restore_gbps']} "
          f"({c3['restore_frac']}), {c3['blocked_ms']} | {c4['worlds']} ({c4['gb']} GB): {c4['take_gbps']} ({c4['take_frac']}), {c4['restore_gbps']} "
          f"({c4['restore_frac']}) | {c5['gbps']} ({c5['gb']} GB) | {rc_txt} | {c5['read_object_1Mrows_100MiB']['ms']} ms |")
