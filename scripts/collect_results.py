"""Collect the results table for profiles/RESULTS.md: every BASELINE.json graph shape, our GPU variant vs our
serial (v1-equivalent) CPU variant vs the reference-style naive GPU loop, same queries.  Run under gpurun."""
import json, sys, time
import numpy as np, torch
sys.path.insert(0, ".")
import bidirectional_bfs_b200 as bb
sys.path.insert(0, ".")
import bench

out = {}
configs = sys.argv[1:] or ["er1k", "er10k", "er50k", "er100k", "er1m", "rmat20", "rmat10m", "grid1024", "grid16m"]
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for cfg in configs:
    spec, desc = bench.make_spec(bb, cfg, 1), bench.spec_string(cfg, 1)[1]
    n = int(spec.n)
    t0 = time.time(); g = bb.Graph.generate(spec, device="cuda:0"); torch.cuda.synchronize(); build_s = time.time() - t0
    solver = bb.GpuSolver(g)
    qs = [(a, b, L) for a, b, L, _, _ in bench.prepare(cfg, 1, 12)["queries"]]
    rec = {"desc": desc, "n": n, "nnz": g.nnz, "build_s": build_s, "queries": len(qs)}
    ms, pms, lev, edges = [], [], [], []
    for s, t, L in qs:
        best = None
        for _ in range(3):
            flush.fill_(1)
            r = solver.query(s, t, timing="events")
            assert r.length == L
            best = r if best is None or r.search_ms < best.search_ms else best
        ms.append(best.search_ms); pms.append(best.path_ms); lev.append(best.levels); edges.append(best.edges_scanned)
    rec.update(gpu_ms_mean=float(np.mean(ms)), gpu_ms_median=float(np.median(ms)), gpu_q0_ms=ms[0], path_ms_mean=float(np.mean(pms)),
               levels_mean=float(np.mean(lev)), len_q0=qs[0][2], own_edges_mean=float(np.mean(edges)),
               us_per_level=float(np.sum(ms) * 1e3 / np.sum(lev)))
    if n <= 20_000_000:
        gc = g.to("cpu")
        ser = bb.SerialSolver(gc, policy=0)
        sms, sed = [], []
        for s, t, L in qs:
            rr = min((ser.query(s, t, want_path=False) for _ in range(3)), key=lambda x: x.search_ms)
            assert rr.length == L
            sms.append(rr.search_ms); sed.append(rr.edges_scanned)
        rec.update(serial_ms_mean=float(np.mean(sms)), serial_q0_ms=sms[0], v1_edges_mean=float(np.mean(sed)),
                   gteps_v1_numerator=float(np.sum(sed) / (np.sum(ms) * 1e-3) / 1e9), serial_gteps=float(np.sum(sed) / (np.sum(sms) * 1e-3) / 1e9))
        del gc
    if n <= 2_000_000:
        nv = bb.NaiveGpuSolver(g)
        nms = []
        for s, t, L in qs[:4]:
            rr = nv.query(s, t); rr = nv.query(s, t)
            nms.append(rr.search_ms)
        rec["naive_gpu_ms_mean"] = float(np.mean(nms))
    solver.close()
    for P in (2, 8):
        if n > 20_000_000: break
        gm = bb.Graph.generate(spec, device="cuda:0", nranks=P)
        m = bb.MultiGpuSolver(gm)
        vms = [min(m.query(s, t, want_path=False).search_ms for _ in range(2)) for s, t, L in qs[:6]]
        rec[f"vranks{P}_ms_mean_first6"] = float(np.mean(vms))
        m.close(); del gm
    rec["gpu_ms_mean_first6"] = float(np.mean(ms[:6]))
    out[cfg] = rec
    print(cfg, json.dumps(rec), flush=True)
    del g; torch.cuda.empty_cache()
json.dump(out, open("gpurun_out/results_1gpu.json", "w"), indent=1)
