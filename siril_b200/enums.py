"""Enum values of the reference, verbatim in meaning and order.
rejection / normalization: src/core/settings.h:34-53; weightingType: src/stacking/stacking.h:46-52;
return codes: src/stacking/stacking.h:18-24."""
NO_REJEC, PERCENTILE, SIGMA, MAD, SIGMEDIAN, WINSORIZED, LINEARFIT, GESDT = range(8)
NO_NORM, ADDITIVE, MULTIPLICATIVE, ADDITIVE_SCALING, MULTIPLICATIVE_SCALING = range(5)
NO_WEIGHT, NBSTARS_WEIGHT, WFWHM_WEIGHT, NOISE_WEIGHT, NBSTACK_WEIGHT = range(5)
ST_OK, ST_GENERIC_ERROR, ST_SEQUENCE_ERROR, ST_CANCEL, ST_ALLOC_ERROR = 0, -1, -2, -9, -10
DATA_USHORT, DATA_FLOAT = 0, 1

REJECTION_NAMES = {NO_REJEC: "none", PERCENTILE: "percentile", SIGMA: "sigma", MAD: "mad",
                   SIGMEDIAN: "sigmedian", WINSORIZED: "winsorized", LINEARFIT: "linearfit", GESDT: "gesdt"}

__all__ = [n for n in dir() if n.isupper()]
