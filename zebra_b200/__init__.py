"""zebra_b200 — B200-native (sm_100a) T-PPR hot path of Zebra behind the reference's own interface.

Public surface (mirrors LuckyLYM/Zebra utils/util.py):
    tppr_finder      streaming top-k T-PPR finder (device tables + CUDA span kernel)
"""
__all__ = ["tppr_finder"]


def __getattr__(name):
    if name == "tppr_finder":
        from .finder import tppr_finder
        return tppr_finder
    raise AttributeError(name)
