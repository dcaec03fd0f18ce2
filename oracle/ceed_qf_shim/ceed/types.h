/* Stand-in for <ceed/types.h> so that the reference's libCEED Q-FUNCTION headers (src/swe/swe_well_balancing.h, swe_fluxes_ceed.h,
 * swe_roe_flux_ceed.h, swe_sources_ceed.h) compile as plain C where they lie under /root/reference.  TEST INFRASTRUCTURE ONLY.
 * A Q-function is an ordinary function over Q quadrature points; `in` / `out` are arrays of component-major fields. */
#ifndef RDYB_CEED_QF_SHIM_TYPES_H
#define RDYB_CEED_QF_SHIM_TYPES_H
#include <math.h>
typedef double CeedScalar;
typedef int    CeedInt;
#define CEED_QFUNCTION(name) static int name
#define CEED_QFUNCTION_HELPER static inline
#define CEED_Q_VLA Q
#endif
