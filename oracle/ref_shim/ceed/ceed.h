/* stand-in for libCEED's header: opaque handles only (TEST INFRASTRUCTURE) */
#ifndef RDYB_REF_SHIM_CEED_H
#define RDYB_REF_SHIM_CEED_H
typedef struct _shim_Ceed       *Ceed;
typedef struct _shim_CeedOp     *CeedOperator;
typedef struct _shim_CeedVec    *CeedVector;
typedef struct _shim_CeedQFC    *CeedQFunctionContext;
typedef struct _shim_CeedER     *CeedElemRestriction;
typedef double                   CeedScalar;
typedef int                      CeedInt;
typedef int                      CeedMemType;
#endif
