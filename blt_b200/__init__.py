"""blt_b200 -- B200-native (sm_100a) implementation of BLT's lattice-search hot path.

Host-side mirror of the reference's interface for that path:
  closest()     <-> libblt/closest.h:50-54     (search over the LLL-reduced basis)
  infdist()     <-> libblt/glpk_wrapper.h:26-34 (one layer-distance solve)
  check_soln()  <-> libblt/closest.cc:330-340
All of them call the CUDA kernels through the C ABI declared in include/bltcu.h.  There is no CPU fallback: if
libbltcu.so is missing or no CUDA device is usable the calls raise / return BLT_STATUS_ERROR.
"""
from .search import (BLT_CHECK_SAT, BLT_CHECK_UNSAT, BLT_STATUS_ERROR, BLT_STATUS_INPUT_ERROR, BLT_STATUS_LP_FAIL,
                     BLT_STATUS_OK, check_soln, closest, closest_batch, closest_dev, closest_noalloc, fp64_peak, infdist, int32_peak, lib, lll, lll_arbitrary, lll_batch, lll_fixed_host, lll_fp_host, version, yprune)

__all__ = ["closest", "closest_batch", "closest_dev", "closest_noalloc", "infdist", "check_soln", "lib", "lll", "lll_arbitrary", "lll_batch", "lll_fixed_host", "lll_fp_host", "version", "yprune", "fp64_peak", "int32_peak", "BLT_STATUS_OK", "BLT_STATUS_ERROR",
           "BLT_STATUS_LP_FAIL", "BLT_STATUS_INPUT_ERROR", "BLT_CHECK_SAT", "BLT_CHECK_UNSAT"]
