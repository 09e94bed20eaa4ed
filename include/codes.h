/* codes.h -- BLT's return codes.  Values are part of the ABI (reference: libblt/codes.h:5-24; the Haskell binding
 * hard-codes them, src/BLT/Internal.hsc:40-45). */
#ifndef BLT_CODES_H
#define BLT_CODES_H

enum return_codes { BLT_STATUS_OK = 0, BLT_STATUS_ERROR = 1, BLT_STATUS_LP_FAIL = 2, BLT_STATUS_INPUT_ERROR = 3 };
enum sat_codes { BLT_CHECK_SAT = 10, BLT_CHECK_UNSAT = 11, BLT_CHECK_UNKNOWN = 12 };
enum lp_codes { LP_OK = 20, LP_ERROR = 21 };

#endif
