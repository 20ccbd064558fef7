"""deepearth_b200 -- B200 (sm_100a) implementation of DeepEarth's Grid4D hash-grid spacetime encoder.

Public surface mirrors the reference's modules for this path:
    MultiResolutionHashEncoding  <- models/hash_encoding.py
    Grid4DEncoder                <- models/encoders.py
    DeepEarthConfig              <- models/configs.py
    GeospatialConverter, geo2xyz <- geospatial/geo2xyz.py (position path)
All compute goes through libde_grid4d.so (include/de_grid4d.h); there is no CPU fallback.
"""
from .configs import DeepEarthConfig, ModalityConfig, TransformerConfig
from .encoders import Grid4DEncoder
from .geo2xyz import BoundingBox, GeospatialConverter, geo2xyz
from .hash_encoding import MultiResolutionHashEncoding
from .st_embedding import SpatialTemporalEmbedding

__all__ = ["SpatialTemporalEmbedding", "DeepEarthConfig", "ModalityConfig", "TransformerConfig", "Grid4DEncoder", "BoundingBox",
           "GeospatialConverter", "geo2xyz", "MultiResolutionHashEncoding"]
__version__ = "0.1.0"
