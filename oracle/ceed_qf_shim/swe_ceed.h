/* What CMake generates from include/private/physics_ceed.h.in for PHYSICS = swe (values of include/private/rdycoreimpl.h).
 * TEST INFRASTRUCTURE ONLY. */
#ifndef RDYB_CEED_QF_SHIM_SWE_CEED_H
#define RDYB_CEED_QF_SHIM_SWE_CEED_H
#include <ceed/types.h>
#define MATERIAL_PROPERTY_MANNINGS 0
#define NUM_MATERIAL_PROPERTIES 1
#define MAX_NUM_TRACERS 5
#ifndef Square
#define Square(x) ((x) * (x))
#endif
#ifndef SafeDiv
#define SafeDiv(a, b, c, tiny) ((c) > (tiny) ? (a) / (b) : 0.0)
#endif
#endif
