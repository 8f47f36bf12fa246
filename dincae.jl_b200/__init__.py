"""dincae_b200 -- host side of the B200-native DINCAE hot path.

A thin mirror of the reference's operator interface (DINCAE.reconstruct,
DINCAE.reconstruct_points, NCData/DataBatches, genmodel/StepModel, train_init/train_epoch!,
interpnd/interp_adjn) over the C-ABI library libdincae_b200.so.  All compute happens in the
CUDA library; there is no CPU fallback: every call raises if the library or a GPU is missing.
"""
from . import lib            # noqa: F401  ctypes declarations of include/dincae_b200.h
from . import api            # noqa: F401  reference-shaped host API
from .api import (Context, Model, GriddedDataset, PointDataset, ModelConfig,
                  interpnd, interp_adjn, reconstruct, reconstruct_points)
