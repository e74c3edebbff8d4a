#!/usr/bin/env python3
"""Markdown tables of one bench.py JSON line (DESIGN.md section 6 is generated from the committed profiles/*.json)."""
import json
import sys


def main(path):
    d = json.loads(open(path).read().strip().splitlines()[-1])
    c = d.get("configs", {})
    rows = [("humanoid3d walk, 4 096 envs (configs[1], headline)", d["ms_per_step"] * 1e3, d["value"], d["roofline"]["frac"], d["roofline"].get("traffic"))]
    names = [("config3_spinkick_dance_b_65536_sharded", "config 3: spinkick + dance_b, 65 536 envs"),
             ("config4_g1_combined_262144_sharded", "config 4: G1 combined, 262 144 envs"),
             ("target_humanoid3d_walk_1M_per_gpu", "humanoid3d walk, 1 M envs per GPU (north-star row)"),
             ("target_g1_combined_1M_per_gpu", "G1 combined, 1 M envs per GPU"),
             ("layout_soa_humanoid3d_walk_1M", "humanoid3d walk, 1 M envs, SoA layout (K1s)"),
             ("layout_soa_g1_combined_1M", "G1 combined, 1 M envs, SoA layout (K1s)")]
    for k, label in names:
        if k in c:
            r = c[k]
            rows.append((label, r["ms_per_step"] * 1e3, r["env_steps_per_s"], r["frac"], r.get("traffic")))
    print("| workload | µs / step | env-steps/s | HBM frac | ncu DRAM bytes / launch |")
    print("|---|---|---|---|---|")
    for label, us, v, f, t in rows:
        print("| %s | %.1f | %.3g | %.3f | %s |" % (label, us, v, f, "%.1f MB" % (t / 1e6) if t else "—"))
    e = d.get("e2e")
    if e:
        print("| e2e `DPVecEnv.step`, 4 096 envs, pageable numpy in / caller-owned arrays out | %.0f | %.3g | — | %.2f MB in + %.2f MB out over PCIe; in place: %.0f µs, %.3g |"
              % (e["ms_per_step"] * 1e3, e["value"], e["h2d_bytes_per_step"] / 1e6, e["d2h_bytes_per_step"] / 1e6,
                 e["pinned"]["ms_per_step"] * 1e3, e["pinned"]["value"]))
    e4 = c.get("config4_g1_combined_262144_sharded", {}).get("e2e")
    if e4:
        print("| e2e, config 4 (262 144 G1 combined envs), pageable / in place | %.0f / %.0f | %.3g / %.3g | — | %.1f MB in + %.1f MB out |"
              % (e4["ms_per_step"] * 1e3, e4["pinned"]["ms_per_step"] * 1e3, e4["value"], e4["pinned"]["value"],
                 e4["h2d_bytes_per_step"] / 1e6, e4["d2h_bytes_per_step"] / 1e6))
    cb = d.get("cpu_baseline")
    if cb:
        sp = cb.get("single_process", {})
        print("| CPU arm (%s, %d cores; %s) | — | %.3g | — | single-process loop: %.3g env-steps/s (%.2f ms per step) |"
              % (cb["kind"], cb["cores"], cb["sample"][:90] + "…", cb["value"], sp.get("value", 0), sp.get("ms_per_step", 0)))
    sw = c.get("config5_sweep")
    if sw:
        print()
        print("| envs | humanoid3d walk µs | frac | G1 combined µs | frac |")
        print("|---|---|---|---|---|")
        for r in sw:
            h, g = r["humanoid3d_walk"], r["g1_combined"]
            print("| %s | %.1f | %.3f | %.1f | %.3f |" % (format(r["envs_per_gpu"], ",").replace(",", " "), h["kernel_ms"] * 1e3, h["frac"], g["kernel_ms"] * 1e3, g["frac"]))


if __name__ == "__main__":
    main(sys.argv[1])
