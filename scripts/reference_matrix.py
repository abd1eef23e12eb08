"""The reference's own benchmark matrix (benchmark_test.sh: 4 variants x {1k,10k,50k,100k}, query 0 -> n-1) on this box:
the unmodified reference binaries (baseline/_ref/bin, v2/v4 through the MPI stand-in with np 4) next to our variants.
Writes gpurun_out/reference_matrix.md (+ our benchmark_results.csv / benchmark_table.txt)."""
import os, re, subprocess, sys, statistics
from pathlib import Path
sys.path.insert(0, ".")
import bidirectional_bfs_b200 as bb
from bidirectional_bfs_b200 import bench_harness

ROOT = Path(".").resolve()
out = ROOT / "gpurun_out"; out.mkdir(exist_ok=True)
gdir = Path("/tmp/ref_matrix_graphs"); gdir.mkdir(exist_ok=True)
names = ["1k", "10k", "50k", "100k"]
paths = []
for nm in names:
    spec = bb.preset_spec(nm, seed=1)
    p = gdir / f"{nm}.bin"
    bb.io.write_bin(p, int(spec.n), bb.graph.edge_list(spec, "cpu"))
    paths.append(p)
rows, _table = bench_harness.run([str(p) for p in paths], ["serial", "cpu-ranks", "naive-gpu", "gpu", "multi-gpu"], repeats=7, ranks=4, outdir=str(out))
ours = {}
for r in rows:
    ours[(r[0], r[1])] = float(r[2]) * 1e3          # seconds -> ms
refbin = ROOT / "baseline" / "_ref" / "bin"
mpirun = str(ROOT / "baseline" / "mpi_shim" / "mpirun")
env = dict(os.environ, MPISHIM_GPU_PER_RANK="0")
pat = [re.compile(r"took ([0-9.eE+-]+) seconds"), re.compile(r"\[Time\][^=]*=\s*([0-9.eE+-]+)\s*s"), re.compile(r"gpu bfs time = ([0-9.eE+-]+) ms")]
ref = {}
for nm, p in zip(names, paths):
    n = int(bb.preset_spec(nm, seed=1).n)
    for label, cmd in (("v1 as shipped", [str(refbin / "v1_as_shipped")]), ("v1 -O3", [str(refbin / "v1_O3")]),
                       ("v2 MPI np 4", [mpirun, "-n", "4", str(refbin / "v2")]), ("v3 CUDA", [str(refbin / "v3")]),
                       ("v4 MPI+CUDA np 4", [mpirun, "-n", "4", str(refbin / "v4")])):
        if not Path(cmd[-1]).exists() and not Path(cmd[0]).exists():
            continue
        vals = []
        for _ in range(3):
            o = subprocess.run(cmd + [str(p), "0", str(n - 1)], capture_output=True, text=True, env=env, timeout=600).stdout
            for k, rx in enumerate(pat):
                m = rx.search(o)
                if m:
                    vals.append(float(m.group(1)) * (1.0 if k == 2 else 1e3))
                    break
        if vals:
            ref[(label, nm)] = min(vals)
lines = ["# The reference's benchmark matrix on the B200 box (ms per query 0 -> n-1, search only; reference: best of 3 runs of the",
         "# unmodified binaries, v2 / v4 through baseline/mpi_shim with np 4; ours: median of 7, `python -m bidirectional_bfs_b200 bench`)", "",
         "| variant | " + " | ".join(names) + " |", "|---|" + "---|" * len(names)]
for label in ("v1 as shipped", "v1 -O3", "v2 MPI np 4", "v3 CUDA", "v4 MPI+CUDA np 4"):
    lines.append(f"| reference {label} | " + " | ".join(f"{ref[(label, nm)]:.4f}" if (label, nm) in ref else "-" for nm in names) + " |")
keys = sorted({k[0] for k in ours})
for v in keys:
    lines.append(f"| ours {v} | " + " | ".join(f"{ours[(v, nm + '.bin')]:.4f}" if (v, nm + ".bin") in ours else (f"{ours[(v, nm)]:.4f}" if (v, nm) in ours else "-") for nm in names) + " |")
(out / "reference_matrix.md").write_text("\n".join(lines) + "\n")
print("\n".join(lines))
