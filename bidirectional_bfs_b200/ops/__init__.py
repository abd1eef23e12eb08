"""Device/host operators: generators, CSR builder, scan, search context."""
from .gen import er_spec, rmat_spec, grid_spec, preset_spec, generate_edges, count_edges  # noqa: F401
from .csr import build_csr, exclusive_scan_u32  # noqa: F401
