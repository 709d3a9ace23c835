"""vaeclust -- B200-native (sm_100a) VAE training step and latent-clustering evaluation.

Drop-in host layer for the hot path of NawrozHaseen/VAE-for-Hybrid-Music-Clustering (scripts 06/12/19 training,
07/13/16/20 clustering) over the C-ABI of libvaeclust.so (include/vaeclust.h).  Importing this package never
touches the GPU; the shared library is loaded on first use and its absence is an error, not a fallback.
"""
from ._lib import HEADER, LIB_PATH, VaeclustError, lib, parse_header  # noqa: F401

__version__ = "0.1.0"
