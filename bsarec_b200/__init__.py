"""bsarec_b200 — B200-native (sm_100a) drop-in for BSARec's training / full-ranking-eval hot path.

``MODEL_DICT["bsarec"]`` mirrors ``src/model/__init__.py:10-11`` of the reference.
"""
from .model import BSARecModel, SASRecModel, DuoRecModel, FMLPRecModel
from .trainer import Trainer
from .optim import FusedAdam
from .data import SequenceStore, DeviceBatchLoader, HitRankAccumulator
from .graphs import GraphedTrainStep, GraphedTopk

MODEL_DICT = {"bsarec": BSARecModel, "sasrec": SASRecModel, "duorec": DuoRecModel, "fmlprec": FMLPRecModel}

__all__ = ["BSARecModel", "SASRecModel", "DuoRecModel", "FMLPRecModel", "Trainer", "FusedAdam", "MODEL_DICT", "SequenceStore", "DeviceBatchLoader",
           "HitRankAccumulator", "GraphedTrainStep", "GraphedTopk"]
