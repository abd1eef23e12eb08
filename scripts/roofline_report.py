"""Per-level roofline report (run under gpurun): for each traced level the algorithmic bytes over HBM and over NVLink
against the measured HBM copy bandwidth (MEASURED_PEAKS.json) and NVLink 5 at 900 GB/s per direction:

    fraction = max(hbm_bytes / HBM_peak, nvlink_bytes / 900 GB/s) / t_level          (BASELINE.json `metric`)

    python scripts/roofline_report.py <config> [flags] [ranks]        ranks > 1: P ranks (virtual ranks on this GPU — the
                                                                       protocol and the byte counts are those of P GPUs)
Bytes per level (r = (P-1)/P = share of vertices owned by another rank; records = 32-byte vertex records):
  paired / wide rounds of the front kernel (mode & 16): per inspected entry one claim (8 B visited word, rank 0's HBM) and, in
    narrow rounds, the candidate's record (32 B; from the peer's HBM over NVLink for the share r unless rank 0 holds a record
    replica: --replica 1); per frontier vertex of a wide round its log entry + record (4 + 32 B); per discovery 8 B log + parent.
  fused levels of the general kernel: per frontier vertex 4 + 16 B (queue, row offsets), per entry 4 B column + 8 B claim on the
    OWNER (share r over NVLink, 8 B each way), per discovery 8 B entry + 4 B parent (share r over NVLink), per level and peer
    64 B of exchange words."""
import json, sys
import numpy as np, torch
sys.path.insert(0, ".")
import bidirectional_bfs_b200 as bb
import bench
cfg = sys.argv[1]; flags = int(sys.argv[2]) if len(sys.argv) > 2 else 0; P = int(sys.argv[3]) if len(sys.argv) > 3 else 1
replica = "--replica" in sys.argv
try:
    hbm = json.load(open("MEASURED_PEAKS.json"))["hbm_gbs"]; src = "measured"
except Exception:
    hbm = 6650.0; src = "fallback"
NVL = 900.0
spec, desc = bench.make_spec(bb, cfg, 1), bench.spec_string(cfg, 1)[1]
g = bb.Graph.generate(spec, device="cuda:0", nranks=P)
s = bb.GpuSolver(g) if P == 1 else bb.MultiGpuSolver(g)
ctx = s.ctx if P == 1 else s.ctxs[0]
qs = [(a, b, L) for a, b, L, _, _ in bench.prepare(cfg, 1, 3)["queries"]]
r_ = (P - 1) / P
print(f"# {cfg}: {desc}; ranks={P}{' (record replica on rank 0)' if replica else ''}; HBM peak {hbm} GB/s ({src}), NVLink {NVL} GB/s per direction; flags={flags}")
for a, b, L in qs:
    for _ in range(2): r = s.query(a, b, flags=1 | flags)
    tr = ctx.trace()
    print(f"query {a}->{b} len={L} levels={r.levels} (paired {r.levels_paired}) search_ms={r.search_ms:.4f}")
    prev = None; rows = []; tot_h = tot_n = 0
    for i, t in enumerate(tr):
        dt = (t["t_ns"] - prev) * 1e-9 if prev else None
        prev = t["t_ns"]
        f, e, d, m = t["frontier"], t["edges"], t["discovered"], t["mode"]
        if m & 16:
            rec = e * 32 if not (m & 32) else f * 36
            h = e * 8 + d * 8 + (rec if (P == 1 or replica) else rec * (1 - r_))
            n = 0 if (P == 1 or replica) else rec * r_
        else:
            h = f * 20 + e * 4 + e * 8 * (1 - r_) + d * 12 * (1 - r_)
            n = e * 16 * r_ + d * 12 * r_ + (P - 1) * 64
        tot_h += h; tot_n += n
        if dt and dt > 0:
            bound = max(h / (hbm * 1e9), n / (NVL * 1e9))
            rows.append((i, t["side"], m, f, e, d, dt * 1e6, h / 1e3, n / 1e3, bound / dt))
    show = rows if len(rows) <= 12 else rows[:4] + sorted(rows[4:], key=lambda x: -x[9])[:6]
    for x in show:
        print("   L%-5d side=%d mode=%-2d frontier=%-8d edges=%-10d disc=%-8d dt=%8.2f us  HBM %9.1f KB  NVLink %9.1f KB  fraction of roofline %.5f" % x)
    tq = r.search_ms * 1e-3
    print(f"   whole query: HBM {tot_h/1e3:.1f} KB, NVLink {tot_n/1e3:.1f} KB -> {max(tot_h/(hbm*1e9), tot_n/(NVL*1e9))/tq:.5f} of the roofline; {r.edges_scanned / tq / 1e9:.4f} GTEPS (own edges)")
