"""zebra_b200 — B200-native (sm_100a) T-PPR hot path of Zebra behind the reference's own interface.

Public surface (mirrors LuckyLYM/Zebra utils/util.py):
    tppr_finder      streaming top-k T-PPR finder (device tables + CUDA span kernel)
    NeighborFinder, get_neighbor_finder
                     static adjacency + pruned power-iteration top-k query (CUDA)
"""
__all__ = ["tppr_finder", "NeighborFinder", "get_neighbor_finder"]


def __getattr__(name):
    if name == "tppr_finder":
        from .finder import tppr_finder
        return tppr_finder
    if name in ("NeighborFinder", "get_neighbor_finder"):
        from . import neighbor_finder
        return getattr(neighbor_finder, name)
    raise AttributeError(name)
