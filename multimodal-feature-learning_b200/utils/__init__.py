# mirrors the parts of the reference that feed the hot path directly: utils/dam.py and the
# encoders' get_reference_points
from .dam import attn_map_to_flat_grid, compute_corr, idx_to_flat_grid
from .encoder_inputs import prepare_encoder_inputs
from .reference_points import get_reference_points
from .sparse_tokens import gather_tokens, scatter_tokens

__all__ = ["attn_map_to_flat_grid", "compute_corr", "idx_to_flat_grid", "get_reference_points",
           "gather_tokens", "scatter_tokens", "prepare_encoder_inputs"]
