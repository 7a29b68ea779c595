"""continuousnf.jl_b200 — B200-native batched augmented-dynamics solve of
ContinuousNormalizingFlows.jl behind the reference's own operator interface.

The directory name contains a dot, so import it through `cnfb200_loader.load_package()` (repo
root) or importlib; the module registers itself as `cnfb200`.
"""
from . import _cabi, adapters, parallel, workloads  # noqa: F401
from .adapters import CondICNFDist, CondICNFModel, ICNFDist, ICNFModel, Lion, fit, transform  # noqa: F401
from .api import (  # noqa: F401
    B200Libs, Chain, CondFFJORD, CondPlanar, CondRNODE, Dense, Planar, PlanarLayer, DIJacVecMatrixMode, DIJacVecVectorMode,
    DIVecJacMatrixMode, DIVecJacVectorMode, FFJORD, ICNF, RNODE, TestMode, TrainMode, augmented_f, augmented_f_static,
    construct, generate, inference, last_stats, loss, loss_and_grad, param_offsets, set_option, setup,
    steer_tspan, seed_device_rng, last_eps, fit_epochs, steer_sequence,
)
