"""Host mirror of MCMC::Analysis::CovarianceMatrix (MCMCpp/Analysis/CovarianceMatrix.h) over the C-ABI.

Same surface as the reference class: construct with (numParams, numWalkers), `calculateCovar(chain, sliceInterval)`
where the reference takes a ChainStepIterator range the mirror takes what that range walks -- the contiguous rows
[steps][walkers][params] returned by `EnsembleSampler.getChain()` -- then `getCovarianceMatrixElement(row, col)` /
`getCorrelationMatrixElement(row, col)`.  The sums run on the GPU (mcmcpp_b200/csrc/analysis.cu); there is no CPU path.
"""
import ctypes as C

import numpy as np

from . import _capi
from ._capi import check


class CovarianceMatrix:
    def __init__(self, numParams, numWalkers, device=0):
        if numParams <= 0 or numWalkers <= 0:          # the reference asserts (CovarianceMatrix.h:103-104)
            raise _capi.McmcppError(_capi.ERR_ARG, "numParams and numWalkers must be positive")
        self.pCount, self.wCount, self.device = int(numParams), int(numWalkers), int(device)
        self.means = np.zeros(self.pCount)
        self.covarMat = np.zeros((self.pCount, self.pCount))
        self.corrMat = np.zeros((self.pCount, self.pCount))
        self.points = 0

    def calculateCovar(self, chain, sliceInterval=1):
        chain = np.ascontiguousarray(chain, dtype=np.float64)
        if chain.ndim != 3 or chain.shape[1:] != (self.wCount, self.pCount):
            raise ValueError("chain must be [steps][%d][%d]" % (self.wCount, self.pCount))
        n = C.c_longlong()
        check(_capi.lib().mcmcpp_covariance(chain.ctypes.data, chain.shape[0], self.wCount, self.pCount, int(sliceInterval),
                                            self.device, self.means.ctypes.data, self.covarMat.ctypes.data,
                                            self.corrMat.ctypes.data, C.byref(n)))
        self.points = n.value

    def getCovarianceMatrixElement(self, row, col):
        return float(self.covarMat[row, col])

    def getCorrelationMatrixElement(self, row, col):
        return float(self.corrMat[row, col])


class AutoCorrCalc:
    """MCMC::Analysis::AutoCorrCalc (MCMCpp/Analysis/AutoCorrCalc.h:77-89): integrated autocorrelation time of every
    parameter.  `calcAutoCorrTimes(chain, numSamples, numWalkersToUse=0)` takes the rows the reference's iterator range
    walks; only all-walker use is offered (the reference's subset is drawn from a random_device-seeded engine,
    AutoCorrCalc.h:134, and is not reproducible).  `referenceBufferReuse=True` (default) computes what the reference
    computes, including its accumulation of each chain onto the previous walker's autocovariance
    (AutoCorrCalc.h:228-232); False gives the clean definition."""

    def __init__(self, numParams, numWalkers, device=0, referenceBufferReuse=True):
        if numParams <= 0 or numWalkers <= 0:          # the reference asserts (AutoCorrCalc.h:136-137)
            raise _capi.McmcppError(_capi.ERR_ARG, "numParams and numWalkers must be positive")
        self.paramCount, self.walkerCount, self.device = int(numParams), int(numWalkers), int(device)
        self.windowScaling = 4
        self.referenceBufferReuse = bool(referenceBufferReuse)
        self.acorrTimeList = np.zeros(self.paramCount)

    def setAutoCorrScaleFactor(self, scaleFactor=5):
        self.windowScaling = int(scaleFactor)

    def calcAutoCorrTimes(self, chain, numSamples=None, numWalkersToUse=0):
        chain = np.ascontiguousarray(chain, dtype=np.float64)
        if chain.ndim != 3 or chain.shape[1:] != (self.walkerCount, self.paramCount):
            raise ValueError("chain must be [steps][%d][%d]" % (self.walkerCount, self.paramCount))
        if numWalkersToUse not in (0, self.walkerCount):
            raise _capi.McmcppError(_capi.ERR_UNSUPPORTED, "only all-walker autocorrelation times are offered")
        n = chain.shape[0] if numSamples is None else int(numSamples)
        check(_capi.lib().mcmcpp_autocorr_times(chain.ctypes.data, n, self.walkerCount, self.paramCount, self.windowScaling,
                                                int(self.referenceBufferReuse), self.device, self.acorrTimeList.ctypes.data))

    def retrieveAutoCorrelationTime(self, paramIndex):
        return float(self.acorrTimeList[paramIndex])
