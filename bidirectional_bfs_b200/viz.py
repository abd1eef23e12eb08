"""Headless visualisation: draw the graph with the shortest path highlighted, write a PNG.

Counterpart of visual/m3.py (networkx Kamada-Kawai layout, non-path edges invisible, path edges
red width 5, node labels, interactive ``plt.show()``).  There is no display on the target boxes, so
the figure goes to a file: with matplotlib's Agg backend when matplotlib is importable, otherwise
with a small PIL rasteriser (same layout and styling).  The interactive window is unavailable on
this box by design.
"""
from __future__ import annotations

import argparse
import json
import sys


def layout(n, edges, seed=0):
    import networkx as nx
    G = nx.Graph()
    G.add_nodes_from(range(n))
    G.add_edges_from((int(u), int(v)) for u, v in edges)
    try:
        pos = nx.kamada_kawai_layout(G) if n <= 1500 else nx.spring_layout(G, seed=seed, iterations=30)
    except Exception:  # disconnected graphs can upset kamada-kawai on some versions
        pos = nx.spring_layout(G, seed=seed)
    return G, pos


def draw(n, edges, path, out, size=1400, show_all_edges=False, labels=None):
    """Render to `out` (PNG).  Returns the backend used."""
    G, pos = layout(n, edges)
    path_edges = list(zip(path, path[1:])) if path else []
    labels = (n <= 200) if labels is None else labels
    try:
        import matplotlib
        matplotlib.use("Agg")
        import matplotlib.pyplot as plt
        import networkx as nx
        plt.figure(figsize=(size / 100, size / 100))
        nx.draw_networkx_nodes(G, pos, node_size=300, node_color="lightblue")
        nx.draw_networkx_edges(G, pos, edgelist=G.edges(), width=1, alpha=0.15 if show_all_edges else 0.0)
        nx.draw_networkx_edges(G, pos, edgelist=path_edges, edge_color="red", width=5)
        if labels:
            nx.draw_networkx_labels(G, pos, font_size=8)
        plt.axis("off")
        plt.savefig(out, dpi=100)
        plt.close()
        return "matplotlib-agg"
    except ImportError:
        pass
    from PIL import Image, ImageDraw
    img = Image.new("RGB", (size, size), "white")
    d = ImageDraw.Draw(img)
    xs = [p[0] for p in pos.values()]
    ys = [p[1] for p in pos.values()]
    x0, x1, y0, y1 = min(xs), max(xs), min(ys), max(ys)
    m = 40

    def px(v):
        x, y = pos[v]
        return (m + (x - x0) / (x1 - x0 + 1e-12) * (size - 2 * m), m + (1 - (y - y0) / (y1 - y0 + 1e-12)) * (size - 2 * m))
    if show_all_edges:
        for u, v in G.edges():
            d.line([px(u), px(v)], fill=(215, 215, 215), width=1)
    for u, v in path_edges:
        d.line([px(u), px(v)], fill=(255, 0, 0), width=5)
    r = 9 if n <= 200 else 3
    on_path = set(path or [])
    for v in G.nodes():
        x, y = px(v)
        d.ellipse([x - r, y - r, x + r, y + r], fill=(255, 160, 160) if v in on_path else (173, 216, 230), outline=(90, 90, 90))
        if labels or v in on_path:
            d.text((x + r + 1, y - r), str(v), fill=(0, 0, 0))
    img.save(out)
    return "pil"


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="bibfs viz")
    ap.add_argument("graph", help=".bin edge list (or .json graph)")
    ap.add_argument("--path-json", default=None, help="path oracle JSON (default: <stem>.json next to the graph, else computed)")
    ap.add_argument("--src", type=int, default=0)
    ap.add_argument("--dst", type=int, default=None)
    ap.add_argument("--out", default="graph.png")
    ap.add_argument("--all-edges", action="store_true")
    a = ap.parse_args(argv)
    import os
    import bidirectional_bfs_b200 as bb
    n, edges = bb.io.read_json_graph(a.graph) if a.graph.endswith(".json") else bb.io.read_bin(a.graph)
    pj = a.path_json or str(bb.io.oracle_path_for(a.graph))
    if os.path.exists(pj) and (a.path_json or a.dst is None):
        path = bb.io.read_oracle(pj)["nodes"]                     # visual/m3.py:34-36
    else:
        dst = a.dst if a.dst is not None else n - 1
        path = bb.bibfs(bb.Graph.from_edges(n, edges), a.src, dst, variant="serial").path or []
    backend = draw(n, edges.tolist(), path, a.out, show_all_edges=a.all_edges)
    print(f"wrote {a.out} ({backend}); path has {max(0, len(path) - 1)} hops")
    return 0


if __name__ == "__main__":
    sys.exit(main())
