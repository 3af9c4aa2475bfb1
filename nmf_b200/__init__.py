"""nmf_b200 -- B200-native (sm_100a) NMF solvers behind the API of raleng/nmf.

    from nmf_b200 import NMF
    nmf = NMF(data, factors)
    nmf.factorize(method='mur', distance_type='eu')
"""
from .nmf import NMF  # noqa: F401

__all__ = ["NMF"]
