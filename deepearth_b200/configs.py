"""Configuration dataclasses -- field-compatible with the reference's models/configs.py:8-74.

``DeepEarthConfig`` keeps every reference field, default and ``__post_init__`` rule
(resolutions = 2^4.. / 2^2..), and adds opt-in fields for the B200 encoder whose defaults
reproduce the reference behaviour ("Mode R": nearest-cell lookup on an xyz grid + a t grid).
A reference ``DeepEarthConfig`` instance can be passed to ``Grid4DEncoder`` as is -- the extra
fields are read with ``getattr(config, name, default)``.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Any, Dict, List, Optional


@dataclass
class TransformerConfig:
    hidden_dim: int = 768
    n_heads: int = 12
    n_layers: int = 12
    mlp_ratio: float = 4.0
    dropout: float = 0.1
    attention_dropout: float = 0.1
    layer_norm_eps: float = 1e-6


@dataclass
class ModalityConfig:
    name: str
    encoding_type: str
    input_type: str
    column_name: Optional[str] = None
    column_names: Optional[List[str]] = None
    embed_dim: Optional[int] = None
    custom_encoder_config: Optional[Dict[str, Any]] = None


@dataclass
class DeepEarthConfig:
    # --- reference fields (models/configs.py:32-53) -------------------------------------------
    spatial_coordinate_system: str = "cartesian"
    spatial_resolutions: List[int] = None
    temporal_resolutions: List[int] = None
    n_spatial_levels: int = 16
    n_temporal_levels: int = 8
    n_features_per_level: int = 2
    hash_table_size: int = 2 ** 19
    hidden_dim: int = 768
    n_heads: int = 12
    n_layers: int = 12
    modality_configs: Dict[str, ModalityConfig] = None
    modality_encoder_config: TransformerConfig = None
    cross_modal_fusion_config: TransformerConfig = None
    # --- B200 opt-in fields (defaults = reference behaviour) ----------------------------------
    grid4d_mode: str = "reference"          # "reference" (xyz + t, nearest) | "decomposed" (xyz/xyt/yzt/xzt)
    grid4d_interpolation: Optional[str] = None   # None -> "nearest" in reference mode, "linear" in decomposed
    grid4d_table_dtype: str = "float32"     # "float32" | "float16" working tables (fp32 masters kept)
    grid4d_table_scale: Optional[float] = None   # power-of-two scale of fp16 tables (None -> auto)
    grid4d_dense_fallback: Optional[bool] = None  # None -> on in decomposed mode
    grid4d_fused_head: Optional[bool] = None  # projection on tcgen05 kernels (bf16 operands, fp32 accumulate).  None = auto:
                                              # on for the widths the reference uses (hidden_dim a multiple of 128: 768,
                                              # 2048 -> csrc/proj_kernels.cu), off otherwise; True also takes the 64-wide
                                              # one-kernel head (csrc/mlp_head.cu); False = the fp32 nn.Sequential
    grid4d_flat_tables: bool = False        # one (L, T, F) parameter per grid instead of L (T, F) parameters; state dicts
                                            # keep the reference's keys hash_tables.{l}.weight
    spacetime_resolutions: List[int] = None  # resolutions of the xyt/yzt/xzt grids (decomposed mode)

    def __post_init__(self):
        if self.spatial_resolutions is None:
            self.spatial_resolutions = [2 ** i for i in range(4, 4 + self.n_spatial_levels)]
        if self.temporal_resolutions is None:
            self.temporal_resolutions = [2 ** i for i in range(2, 2 + self.n_temporal_levels)]
        if self.modality_encoder_config is None:
            self.modality_encoder_config = TransformerConfig(hidden_dim=self.hidden_dim // 2, n_heads=6, n_layers=4)
        if self.cross_modal_fusion_config is None:
            self.cross_modal_fusion_config = TransformerConfig(hidden_dim=self.hidden_dim, n_heads=self.n_heads,
                                                               n_layers=self.n_layers)
        if self.modality_configs is None:
            self.modality_configs = {}
        if self.grid4d_mode not in ("reference", "decomposed"):
            raise ValueError(f"grid4d_mode must be 'reference' or 'decomposed', got {self.grid4d_mode!r}")
