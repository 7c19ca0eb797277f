"""B200-native engine for the hand-ranking / showdown / equity hot path of sabraham/Monte-Carlo.

The directory name (monte-carlo_b200) is not a Python identifier; load it with `mcb200_loader.load()` at the
repo root, which registers it as the module `monte_carlo_b200`.

  engine          ctypes binding of libmcpoker.so (the C ABI of include/mcpoker.h)
  api             numpy wrappers over the ABI
  card            mirror of montecarlo.card            (card.clj)
  hand_evaluator  mirror of montecarlo.hand-evaluator  (hand_evaluator.clj) -- same names, GPU-backed
  gameplay        mirror of the showdown part of montecarlo.gameplay (gameplay.clj:104-133)
  sharding        index-range split across ranks + the one all-reduce of counts
"""
from . import engine, api, card, hand_evaluator, gameplay, sharding  # noqa: F401
