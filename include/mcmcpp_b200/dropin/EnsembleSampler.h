/*
 * include/mcmcpp_b200/dropin/EnsembleSampler.h -- forwarding header named like the reference's MCMCpp/EnsembleSampler.h.
 * Put  -I<repo>/include/mcmcpp_b200/dropin  BEFORE  -I<reference>/MCMCpp  (and -I<repo>/include anywhere): a program written
 * against the reference, e.g. test/sequential/SkewedGaussian/StretchMove/src/main.cpp with its
 *     #include"Movers/StretchMove.h"
 *     #include"EnsembleSampler.h"
 * then gets the B200 sampler and movers (same namespaces, names and signatures) with no source edit, while Chain/, Analysis/
 * and Utility/ still come from the reference (see include/mcmcpp_b200/EnsembleSampler.h: reference-chain back-end).
 */
#ifndef MCMCPP_B200_DROPIN_ENSEMBLESAMPLER_H
#define MCMCPP_B200_DROPIN_ENSEMBLESAMPLER_H
#include "mcmcpp_b200/EnsembleSampler.h"
#endif
