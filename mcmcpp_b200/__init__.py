"""mcmcpp_b200 -- B200-native ensemble-update hot path of jmatta1/MCMCpp.

Host-side mirror of the reference's sampler interface over the C-ABI runtime
(libmcmcpp_b200.so, include/mcmcpp_b200/capi.h).  Importing the package does not load the
library; the first sampler does, and fails loudly if it was not built."""
from .sampler import (EnsembleSampler, ParallelEnsembleSampler, StretchMove, DifferentialEvolution, WalkMove, SkewedGaussianTwoDim, RosenbrockTwoDim,
                      DiagGaussian, DenseGaussian, LinearRegression)
from .analysis import CovarianceMatrix, AutoCorrCalc

__all__ = ["EnsembleSampler", "ParallelEnsembleSampler", "StretchMove", "DifferentialEvolution", "WalkMove", "SkewedGaussianTwoDim", "RosenbrockTwoDim",
           "DiagGaussian", "DenseGaussian", "LinearRegression", "CovarianceMatrix", "AutoCorrCalc"]
