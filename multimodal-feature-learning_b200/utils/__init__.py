# mirrors the part of the reference's utils/ that touches the hot path's data (utils/dam.py)
from .dam import attn_map_to_flat_grid, compute_corr, idx_to_flat_grid

__all__ = ["attn_map_to_flat_grid", "compute_corr", "idx_to_flat_grid"]
