"""Writes profiles/r02_traffic.json: DRAM bytes of one launch per benchmarked kernel, taken from `ncu --set full` captures of
ONE state of the kernel sources, keyed by bench.lib_sha16() (sha256 over csrc/ sources + Makefile + mi_b200.h) so that bench.py
only reports `roofline.traffic` when the capture on file is of the kernels it is timing.

    python profiles/ncu_traffic.py <lib_sha16 | auto> key=report.ncu-rep:n_samples [key=...]

keys used by bench.py: c2_forward, c3_forward_mono, c3_forward_mono_polarized, c4_adjoint.  Also kept per kernel: the issue-slot
utilisation, active lanes per instruction, L2 hit rate and warps per scheduler of the same capture (bench.py `roofline.issue`)."""
import csv
import json
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))


def dram_bytes(rep):
    raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units, vals = rows[0], rows[1], rows[2]
    d = {h: (v, u) for h, u, v in zip(hdr, units, vals)}
    total = 0.0
    for k in ('dram__bytes_read.sum', 'dram__bytes_write.sum'):
        v, u = d[k]
        total += float(v.replace(',', '')) * {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}[u]
    extra = {}
    for k, name in (('smsp__issue_active.avg.pct_of_peak_sustained_active', 'issue_slots_busy_pct'),
                    ('smsp__thread_inst_executed_per_inst_executed.ratio', 'active_lanes_per_instruction'),
                    ('lts__t_sector_hit_rate.pct', 'l2_hit_rate_pct'),
                    ('smsp__warps_active.avg.per_cycle_active', 'warps_per_scheduler')):
        if k in d:
            extra[name] = round(float(d[k][0].replace(',', '')), 2)
    return total, d['Kernel Name'][0], extra


def main():
    sha = sys.argv[1]
    if sha == 'auto':
        sys.path.insert(0, os.path.dirname(HERE))
        import bench
        sha = bench.lib_sha16()
    out = {"lib_sha16": sha, "kernels": {}}
    for arg in sys.argv[2:]:
        key, rest = arg.split('=', 1)
        rep, n = rest.rsplit(':', 1)
        b, name, extra = dram_bytes(rep)
        out["kernels"][key] = {"dram_bytes": b, "n_samples": int(n), "kernel": name, "report": os.path.basename(rep), "issue": extra}
    json.dump(out, open(os.path.join(HERE, 'r02_traffic.json'), 'w'), indent=1)
    print(json.dumps(out, indent=1))


if __name__ == '__main__':
    main()
