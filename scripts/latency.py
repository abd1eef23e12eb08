"""Latency probes: launch floor (src==dst), per-level cost, clocks while running."""
import subprocess, sys, time
import numpy as np, torch
sys.path.insert(0, ".")
import bidirectional_bfs_b200 as bb

def clocks():
    out = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,clocks.max.sm,clocks.mem,power.draw,pstate", "--format=csv,noheader"], capture_output=True, text=True).stdout.strip()
    return out

spec = bb.preset_spec("100k", seed=1)
g = bb.Graph.generate(spec, device="cuda:0")
for mc in (0,):
    s = bb.GpuSolver(g, max_ctas=mc)
    print("max_ctas", mc, "idle clocks:", clocks())
    ms = [s.query(5, 5, want_path=False, timing="events").search_ms for _ in range(300)]
    print("  src==dst (launch floor): min %.4f med %.4f" % (min(ms), np.median(ms)))
    ms = [s.query(0, spec.n - 1, want_path=False, timing="events").search_ms for _ in range(300)]
    print("  0->n-1 14 levels: min %.4f med %.4f" % (min(ms), np.median(ms)), "clocks now:", clocks())
    # keep GPU busy with a big memset between queries
    buf = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    ms = []
    for _ in range(100):
        buf.fill_(1)
        ms.append(s.query(0, spec.n - 1, want_path=False, timing="events").search_ms)
    print("  with 256MB L2 flush before each: min %.4f med %.4f" % (min(ms), np.median(ms)), "clocks:", clocks())
    r = s.query(0, spec.n - 1, flags=1, timing="events")
    tr = s.ctx.trace()
    print("  level dts us:", " ".join("%.1f" % ((tr[i]["t_ns"] - tr[i - 1]["t_ns"]) / 1e3) for i in range(1, len(tr))))
    print("  expand part us:", " ".join("%.1f" % ((tr[i]["t_expand_ns"] - tr[i - 1]["t_ns"]) / 1e3) for i in range(1, len(tr))))
    print("  sync part us:", " ".join("%.1f" % ((tr[i]["t_ns"] - tr[i]["t_expand_ns"]) / 1e3) for i in range(1, len(tr))))
    s.close()
