"""Regenerate profiles/sass_evidence.txt and docs/sass/*.sass from the built library (no GPU needed).
For every kernel: how often the instructions that show what it is built from occur, with one example each;
plus excerpts of the hot sequences (claim batch of the solo loop, TMA chunk fetch, cluster barrier, exchange)."""
import re, subprocess, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
LIB = ROOT / "bidirectional_bfs_b200" / "_native" / "libbibfs_b200.so"
sass = subprocess.run(["cuobjdump", "-sass", str(LIB)], capture_output=True, text=True).stdout
PAT = [("UBLKCP", "TMA 1-D bulk copy (cp.async.bulk)"), ("SYNCS", "mbarrier arrive / try_wait"), ("UCGABAR", "cluster barrier (barrier.cluster)"),
       (r"ATOMG\.E\.OR\.64\.STRONG\.GPU", "claim: atom.or.b64 on a visited word"), (r"ATOMG\.E\.OR\.64\.STRONG\.SYS", "claim on a peer GPU's word (NVLink)"),
       (r"ST[G]?\.E\.64\.STRONG\.SYS", "tagged 8-byte stores to peers (exchange slots, queue entries)"), (r"LD[G]?\.E\.64\.STRONG\.SYS", "relaxed sys polls (slots, tagged entries)"),
       (r"LDG\.E\.STRONG\.GPU", "strong column loads of a claim batch"), ("ATOMS", "shared-memory atomics (cursors, accumulators)"), (r"MAPA|\.CLUSTER|LDS\.E\.U|STS\.E", "DSMEM access (mapa / shared::cluster)"),
       ("CCTL.E.PF2", "prefetch.global.L2 of row offsets"), ("REDUX", "redux.sync"), ("MATCH", "match.any aggregation"), ("VOTE", "ballot / any / all"), ("SHFL", "shuffles"),
       ("NANOSLEEP", "back-off of parked CTAs"), (r"MEMBAR\.ALL\.SYS", "fence.acq_rel.sys (bottom-up pushes, final exchange only)"), (r"LDG\.E\.(64\.)?CONSTANT", "read-only CSR loads"),
       (r"LDG\.E\.128\.STRONG\.GPU", "32-byte vertex records (two 128-bit loads) of the pair tier"), (r"RED\.E\.OR\.STRONG\.SYS|REDG\.E\.OR\.STRONG\.SYS", "hand-over: visited bits to the owners (red.or.sys)"),
       ("POPC", "popc (incl. the zero_after scheduling fence)"), (r"HMMA|UTCMMA|UTCHMMA|QGMMA", "tensor-core instructions")]
out = ["# cuobjdump -sass bidirectional_bfs_b200/_native/libbibfs_b200.so (sm_100a): occurrences of the instructions that show what each",
       "# kernel is built from.  Regenerate with: python scripts/sass_evidence.py.  There are deliberately no tensor-core instructions:",
       "# BFS has no GEMM-shaped work.", ""]
funcs = re.split(r"\n\s*Function : ", sass)[1:]
docs = ROOT / "docs" / "sass"; docs.mkdir(parents=True, exist_ok=True)
for f in funcs:
    name, body = f.split("\n", 1)
    dem = subprocess.run(["c++filt", name.strip()], capture_output=True, text=True).stdout.strip()
    lines = [l for l in body.splitlines() if re.search(r"/\*[0-9a-f]{4,}\*/", l)]
    out.append(f"## {dem}   ({len(lines)} instructions)")
    for pat, what in PAT:
        hits = [l for l in lines if re.search(pat, l)]
        if hits:
            out.append(f"  {len(hits):6d}  {what}")
            out.append("          e.g. " + re.sub(r"\s+", " ", hits[0].split("/* 0x")[0]).strip())
    out.append("")
    # full listing of the kernel (encodings stripped): docs/sass/full/<kernel>.sass
    short = re.sub(r"[^A-Za-z0-9_]+", "_", re.sub(r"\(.*", "", dem.replace("(anonymous namespace)::", "").replace("bibfs::", "").replace("void ", ""))).strip("_")
    full = docs / "full"; full.mkdir(exist_ok=True)
    (full / f"{short}.sass").write_text(f"// {dem}\n" + "\n".join(re.sub(r"\s+/\* 0x[0-9a-f]+ \*/", "", l) for l in lines) + "\n")
    if "k_front" in dem and ("false" in dem or "(bool)0" in dem):
        ld = [i for i, l in enumerate(lines) if "LDG.E.128.STRONG.GPU" in l]
        runs = [i for k, i in enumerate(ld) if k + 11 < len(ld) and ld[k + 11] - i < 60]
        if runs:
            seg = lines[max(0, runs[0] - 30): runs[0] + 140]
            (docs / "k_front_pair_batch.sass").write_text("\n".join(re.sub(r"\s+/\* 0x[0-9a-f]+ \*/", "", l) for l in seg) + "\n")
        cb = [i for i, l in enumerate(lines) if "UCGABAR" in l]
        if cb:
            seg = lines[max(0, cb[-1] - 30): cb[-1] + 30]
            (docs / "k_front_wide_round_cluster_barrier.sass").write_text("\n".join(re.sub(r"\s+/\* 0x[0-9a-f]+ \*/", "", l) for l in seg) + "\n")
    if "k_bibfs" in dem and "false" in dem or "k_bibfs<(bool)0>" in dem:
        # excerpts: first 8-atomic batch with its loads, first UBLKCP neighbourhood, first cluster barrier
        def excerpt(tag, idx, before, after):
            seg = lines[max(0, idx - before): idx + after]
            (docs / f"k_bibfs_single_{tag}.sass").write_text("\n".join(re.sub(r"\s+/\* 0x[0-9a-f]+ \*/", "", l) for l in seg) + "\n")
        at = [i for i, l in enumerate(lines) if "ATOMG.E.OR.64.STRONG.GPU" in l]
        runs = [i for k, i in enumerate(at) if k + 7 < len(at) and at[k + 7] - i < 80]
        if runs: excerpt("claim_batch", runs[-1], 160, 120)
        ub = [i for i, l in enumerate(lines) if "UBLKCP" in l]
        if ub: excerpt("hub_tma_fetch", ub[0], 40, 60)
        cb = [i for i, l in enumerate(lines) if "UCGABAR" in l]
        if cb: excerpt("cluster_barrier", cb[0], 20, 20)
    if "k_bibfs" in dem and ("true" in dem or "k_bibfs<(bool)1>" in dem):
        st = [i for i, l in enumerate(lines) if re.search(r"STG\.E\.64\.STRONG\.SYS", l)]
        quad = [i for k, i in enumerate(st) if k + 3 < len(st) and st[k + 3] - i < 16]
        if quad:
            seg = lines[max(0, quad[0] - 20): quad[0] + 60]
            (docs / "k_bibfs_multi_exchange.sass").write_text("\n".join(re.sub(r"\s+/\* 0x[0-9a-f]+ \*/", "", l) for l in seg) + "\n")
(ROOT / "profiles" / "sass_evidence.txt").write_text("\n".join(out))
print("kernels:", len(funcs), "-> profiles/sass_evidence.txt, docs/sass/")
