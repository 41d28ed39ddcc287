"""hnsw_b200 -- B200-native (sm_100a) drop-in for the data-parallel hot path of TFMV/hnsw.

Host-side mirror of the reference's Go API over the C ABI in include/b200hnsw.h.  No CPU fallback.
"""
from .distance import (CosineDistance, EuclideanDistance, RegisterDistanceFunc, distanceFuncs, distanceFuncToName)
from .encode import LoadSavedGraph, SavedGraph
from .exact import ExactIndex, NewExactIndex
from ._lib import B200Error
from .graph import Graph, LevelRand, MakeNode, NewGraph, NewGraphWithConfig, Node, maxLevel

__all__ = ["B200Error", "LoadSavedGraph", "SavedGraph", "ExactIndex", "NewExactIndex", "Graph", "LevelRand", "Node", "MakeNode", "NewGraph", "NewGraphWithConfig", "maxLevel", "CosineDistance",
           "EuclideanDistance", "RegisterDistanceFunc", "distanceFuncs", "distanceFuncToName"]
